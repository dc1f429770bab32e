"""The C-ABI library loads and exports exactly what include/vfi_b200.h declares; host-only entry points
(vfi_route_decide) are checked against the oracle; device entry points fail loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import vfi_b200  # noqa: F401
from oracle import router_ref
from vfi_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "vfi_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vfi_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(lib_built):
    names = _declared()
    assert len(names) >= 16
    for n in names:
        assert hasattr(lib_built, n), f"{n} declared in include/vfi_b200.h but not exported"
    assert sorted(_abi.SIGNATURES) == names          # the ctypes table binds every declared symbol, and only those


def test_no_torch_types_in_header():
    src = open(os.path.join(ROOT, "include", "vfi_b200.h")).read()
    code = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    assert "torch" not in code.lower() and "at::" not in code and "#include <torch" not in src and 'extern "C"' in src


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_create_fails_without_gpu(lib_built):
    h = C.c_void_p()
    assert lib_built.vfi_create(0, C.byref(h)) == _abi.E_CUDA and not h.value
    from vfi_b200.engine import Engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Engine(0)


def test_route_decide_matches_oracle(lib_built):
    rng = np.random.default_rng(0)
    for _ in range(2000):
        ssim = rng.uniform(0.3, 1.0)
        mean, mx, std, lv = rng.uniform(0, 12), rng.uniform(0, 60), rng.uniform(0, 30), rng.uniform(0, 900)
        route, conf, hp, ps = _abi.route_decide(ssim, mean, mx, std, lv)
        if ssim < 0.65:
            assert route == _abi.ROUTE_SCENE_CHANGE and conf == 1.0 - ssim
            continue
        want_ps = float(np.sqrt(min(std / 20.0, 1.0) * min(lv / 500.0, 1.0)))
        want_model, want_conf = router_ref.decide(want_ps > 0.4, want_ps, mean, mx, 5.0, 25.0)
        assert ps == want_ps and hp == (want_ps > 0.4)
        assert _abi.ROUTE_NAMES[route] == want_model and conf == float(want_conf)


def test_route_decide_threshold_edges(lib_built):
    # strict comparisons, as adaptive_pipeline.py:244,:266,:269
    assert _abi.route_decide(0.65, 1, 1, 0, 0)[0] == _abi.ROUTE_RIFE
    assert _abi.route_decide(np.nextafter(0.65, 0), 1, 1, 0, 0)[0] == _abi.ROUTE_SCENE_CHANGE
    assert _abi.route_decide(0.9, 1, 25.0, 0, 0)[0] == _abi.ROUTE_RIFE
    assert _abi.route_decide(0.9, 1, np.nextafter(25.0, 26), 0, 0)[0] == _abi.ROUTE_VFIMAMBA
    assert _abi.route_decide(0.9, 5.0, 1, 0, 0)[1] == 0.7
    assert _abi.route_decide(0.9, 0.0, 1, 0, 0)[1] == 1.0


def test_product_path_does_not_import_oracle():
    pkg = os.path.join(ROOT, "video-frame-interpolation-a-benchmark-study-and-lessons-from-failed-novel-approache_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                s = open(os.path.join(d, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", s, flags=re.M), f"{f} imports the oracle"
