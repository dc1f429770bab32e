"""Shared fixtures.  `-m "not gpu"` covers the oracle against the golden vectors, the host logic and the
C-ABI surface; `-m gpu` holds the parity tests proper (CUDA path vs oracle, through the C ABI)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
# the v4.25 release weights cannot be fetched here: the tests opt into the seeded random-init IFNet (rife_wrapper.py
# raises FileNotFoundError without this, as the reference does; test_host_logic.py checks that default)
os.environ.setdefault("VFI_B200_RANDOM_WEIGHTS", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU; run with -m gpu")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    return lambda name: np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="session")
def lib_built():
    """The in-tree libvfi_b200.so (built by __graft_entry__.build())."""
    import vfi_b200  # noqa: F401
    from vfi_b200 import _abi
    if not os.path.exists(_abi.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _abi.lib()


@pytest.fixture(scope="session")
def weights():
    import vfi_b200  # noqa: F401
    from vfi_b200 import ifnet_params
    return ifnet_params.random_init(1234)


@pytest.fixture(scope="session")
def oracle_model(weights):
    from oracle import ifnet_ref
    m = ifnet_ref.Model()
    m.flownet.load_state_dict(weights, strict=True)
    return m


@pytest.fixture(scope="session")
def engine(lib_built, weights):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from vfi_b200.engine import Engine
    e = Engine.get(0)
    e.load_weights(weights)
    return e


@pytest.fixture(scope="session")
def farneback_cv2():
    """cv2.calcOpticalFlowFarneback on RGB frames with the reference's parameters (adaptive_pipeline.py:119-128)."""
    import cv2

    def run(f0, f1, levels=3, iterations=3, poly_sigma=1.2):
        g0 = cv2.cvtColor(f0, cv2.COLOR_RGB2GRAY) if f0.ndim == 3 else f0
        g1 = cv2.cvtColor(f1, cv2.COLOR_RGB2GRAY) if f1.ndim == 3 else f1
        return cv2.calcOpticalFlowFarneback(g0, g1, None, 0.5, levels, 15, iterations, 5, poly_sigma, 0)
    return run


# the reference's own fixtures (tests/conftest.py:15-36), seeded
@pytest.fixture
def sample_rgb_frame():
    return np.random.default_rng(0).integers(0, 256, (480, 640, 3), dtype=np.uint8)


@pytest.fixture
def sample_frame_pair(sample_rgb_frame):
    return sample_rgb_frame, np.roll(sample_rgb_frame, 5, axis=1)
