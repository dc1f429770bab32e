"""CUDA Farneback (csrc/farneback.cu) against cv2.calcOpticalFlowFarneback and against the
numpy restatement (oracle/farneback_ref.py, itself pinned against cv2).  Parity with OpenCV's internals is floating
point, so the bar is the flow tolerance VERDICT r01 set (EPE <= 0.05 px mean / 0.5 px p99) -- measured far inside it --
plus identical routing decisions (tests/test_gpu_router.py)."""
import cv2
import numpy as np
import pytest
import torch

import vfi_b200  # noqa: F401
from oracle import farneback_ref
from vfi_b200.synth import gen_clip

pytestmark = pytest.mark.gpu


def _epe(a, b):
    return np.sqrt(((a - b) ** 2).sum(-1))


@pytest.mark.parametrize("H,W,seed", [(135, 240, 1), (203, 317, 3), (360, 640, 4), (480, 640, 6), (1080, 1920, 2)])
def test_flow_vs_cv2(engine, farneback_cv2, H, W, seed):
    clip = gen_clip(2, H, W, seed=seed, cut_every=0, fast=(0, 2))
    flow, st = engine.farneback(clip[:2], clip[1:3], want_stats=True)
    flow, st = flow.cpu().numpy(), st.cpu().numpy()
    for i in range(2):
        ref = farneback_cv2(clip[i], clip[i + 1])
        e = _epe(flow[i], ref)
        assert e.mean() <= 1e-3 and np.percentile(e, 99) <= 1e-2 and e.max() <= 0.05, (e.mean(), e.max())
        mag = np.sqrt(ref[..., 0] ** 2 + ref[..., 1] ** 2)
        assert abs(st[i, 0] - mag.mean()) <= 1e-4 and abs(st[i, 1] - mag.max()) <= 5e-2 and abs(st[i, 2] - mag.std()) <= 1e-4
        assert mag.max() > 3.0


def test_flow_vs_restatement_and_gray_input(engine):
    clip = gen_clip(1, 203, 317, seed=3, cut_every=0, fast=(0, 2))
    g = np.stack([cv2.cvtColor(f, cv2.COLOR_RGB2GRAY) for f in clip])
    want = farneback_ref.farneback(g[0], g[1])
    from_rgb = engine.farneback(clip[:1], clip[1:2])[0].cpu().numpy()
    from_gray = engine.farneback(g[:1], g[1:2])[0].cpu().numpy()
    assert np.array_equal(from_rgb, from_gray)                           # RGB2GRAY on the device = cv2's integers
    assert float(np.abs(from_rgb - want).max()) <= 1e-3


def test_single_level_variant(engine, farneback_cv2):
    """scripts/run_experiments.py:359-362: Farneback(0.5, 1, 15, 1, 5, 1.1) on 480x270 grays, mean |flow| routed on."""
    clip = gen_clip(1, 270, 480, seed=5, cut_every=0, fast=None)
    g = np.stack([cv2.cvtColor(f, cv2.COLOR_RGB2GRAY) for f in clip])
    ref = farneback_cv2(g[0], g[1], levels=1, iterations=1, poly_sigma=1.1)
    st = engine.farneback(g[:1], g[1:2], levels=1, iterations=1, poly_sigma=1.1, want_flow=False, want_stats=True).cpu().numpy()[0]
    assert abs(st[0] - np.sqrt((ref ** 2).sum(-1)).mean()) <= 1e-4


def test_batch_equals_singles_and_tiny_frames(engine, farneback_cv2):
    clip = gen_clip(11, 96, 128, seed=9, cut_every=0, fast=None)        # 11 pairs: two chunks of the 8-pair batching
    both = engine.farneback(clip[:-1], clip[1:]).cpu().numpy()
    for i in (0, 7, 8, 10):
        one = engine.farneback(clip[i:i + 1], clip[i + 1:i + 2])[0].cpu().numpy()
        assert np.array_equal(one, both[i])
    small = gen_clip(1, 40, 56, seed=1, cut_every=0, fast=None)          # under 64 px: zero pyramid levels below full res
    got = engine.farneback(small[:1], small[1:2])[0].cpu().numpy()
    assert _epe(got, farneback_cv2(small[0], small[1])).max() <= 0.05


def test_rejects_unsupported_parameters(engine):
    from vfi_b200 import _abi
    z = np.zeros((1, 64, 64, 3), np.uint8)
    for kw in (dict(poly_n=6), dict(winsize=33), dict(pyr_scale=1.0)):
        with pytest.raises(_abi.VFIError, match="unsupported Farneback"):
            engine.farneback(z, z, **kw)
