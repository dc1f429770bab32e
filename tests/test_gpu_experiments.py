"""Experiment-runner call sites on the GPU against vectors produced by the
reference's own scripts/run_experiments.py (tests/golden/experiments.npz) and against cv2 called directly:
global PSNR bit-identical, global SSIM 1e-12, routing decisions identical, resize / addWeighted / Canny bit-exact,
rife_interpolate (BGR, BORDER_REFLECT pad-128, scale) within the IFNet tolerances."""
import cv2
import numpy as np
import pytest
import torch

import vfi_b200  # noqa: F401
from oracle import experiments_ref as xr
from vfi_b200 import _abi
from vfi_b200.scripts import run_experiments as rx
from vfi_b200.synth import gen_clip, noisy_copy

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def clip(golden):
    g = golden("experiments.npz")
    n, H, W, seed = g["clip"].tolist()
    rgb = gen_clip(n, H, W, seed=seed, cut_every=0, fast=tuple(g["fast"].tolist()))
    return np.ascontiguousarray(rgb[..., ::-1]), np.ascontiguousarray(noisy_copy(rgb, seed=5, amp=4)[..., ::-1]), g


def test_global_psnr_ssim_vs_reference(engine, clip):
    bgr, pred, g = clip
    n = len(bgr) - 1
    ps = [rx.calc_psnr(pred[i], bgr[i]) for i in range(n + 1)] + [rx.calc_psnr(bgr[i], bgr[i + 1]) for i in range(n)]
    ss = [rx.calc_ssim(pred[i], bgr[i]) for i in range(n + 1)] + [rx.calc_ssim(bgr[i], bgr[i + 1]) for i in range(n)]
    assert ps == g["psnr"].tolist()                                         # exact integer SSE -> same float64 expression
    assert np.abs(np.array(ss) - g["ssim"]).max() <= 1e-12
    assert rx.calc_psnr(bgr[0], bgr[0]) == float("inf") and rx.calc_ssim(bgr[0], bgr[0]) == 1.0
    a, b = np.zeros((7, 13, 3), np.uint8), np.full((7, 13, 3), 255, np.uint8)   # odd size: the unaligned tail path
    assert rx.calc_psnr(a, b) == xr.calc_psnr(a, b) == 0.0 and rx.calc_ssim(a, b) == xr.calc_ssim(a, b)


def test_moments_exact(engine):
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (3, 1080, 1920, 3), dtype=np.uint8)
    b = rng.integers(0, 256, (3, 1080, 1920, 3), dtype=np.uint8)
    m = engine.moments(a, b).cpu().numpy()
    for i in range(3):
        x, y = a[i].astype(np.int64).ravel(), b[i].astype(np.int64).ravel()
        assert m[i].tolist() == [int(x.sum()), int(y.sum()), int((x * x).sum()), int((y * y).sum()), int((x * y).sum())]


@pytest.mark.parametrize("H,W,dsize", [(540, 960, (480, 270)), (1080, 1920, (480, 270)), (480, 640, (480, 270)), (271, 333, (480, 270)),
                                      (270, 480, (480, 270)), (1000, 1777, (320, 180))])
def test_resize_gray_bit_exact(engine, H, W, dsize):
    rgb = np.random.default_rng(H).integers(0, 256, (2, H, W, 3), dtype=np.uint8)
    bgr = np.ascontiguousarray(rgb[..., ::-1])
    want = np.stack([cv2.resize(cv2.cvtColor(f, cv2.COLOR_BGR2GRAY), dsize) for f in bgr])
    assert np.array_equal(engine.resize_gray(bgr, dsize, bgr=True).cpu().numpy(), want)
    assert np.array_equal(engine.resize_gray(rgb, dsize).cpu().numpy(), want)
    gray = np.stack([cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in bgr])
    assert np.array_equal(engine.resize_gray(gray, dsize).cpu().numpy(), want)


@pytest.mark.parametrize("t", [0.25, 0.5, 0.75, 0.3, 1.0 / 3.0])
def test_add_weighted_bit_exact(engine, t):
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (2, 270, 481, 3), dtype=np.uint8)
    b = rng.integers(0, 256, (2, 270, 481, 3), dtype=np.uint8)
    assert np.array_equal(engine.add_weighted(a, 1 - t, b, t).cpu().numpy(), cv2.addWeighted(a, 1 - t, b, t, 0))
    flat = np.repeat(np.arange(256, dtype=np.uint8)[:, None], 256, 1)        # every (a, b) byte pair
    assert np.array_equal(engine.add_weighted(flat, 1 - t, flat.T.copy(), t).cpu().numpy(), cv2.addWeighted(flat, 1 - t, flat.T.copy(), t, 0))


@pytest.mark.parametrize("H,W,kind", [(540, 960, "clip"), (1080, 1920, "clip"), (270, 481, "noise"), (97, 65, "noise"), (360, 640, "blobs")])
def test_canny_bit_exact(engine, H, W, kind):
    rng = np.random.default_rng(W)
    if kind == "clip":
        g = np.stack([cv2.cvtColor(f, cv2.COLOR_RGB2GRAY) for f in gen_clip(1, H, W, seed=4, cut_every=0, fast=None)])
    elif kind == "noise":
        g = cv2.GaussianBlur(rng.integers(0, 256, (2, H, W), dtype=np.uint8).transpose(1, 2, 0), (0, 0), 1.5).transpose(2, 0, 1).copy()
    else:
        g = np.zeros((2, H, W), np.uint8)
        for k in range(2):
            for _ in range(40):
                cv2.circle(g[k], (int(rng.integers(0, W)), int(rng.integers(0, H))), int(rng.integers(5, 60)), int(rng.integers(40, 255)), int(rng.integers(1, 9)))
            g[k] = cv2.GaussianBlur(g[k], (0, 0), 1.0)
    edges, dens = engine.canny(g, 100, 200)
    want = np.stack([cv2.Canny(x, 100, 200) for x in g])
    assert np.array_equal(edges.cpu().numpy(), want)
    assert dens.tolist() == [float(np.mean(w > 0)) for w in want]
    if kind != "clip":
        assert (want > 0).mean() > 0.01                                         # the test really has edges (and weak chains)


def test_motion_router_decisions_vs_reference(engine, clip):
    bgr, _, g = clip
    n = len(bgr) - 1
    gold = g["motion"]
    thresh = float(np.sort(gold)[n // 2 - 1] + np.sort(gold)[n // 2]) / 2        # splits the clip: both branches fire
    for i in range(n):
        frames, used, motion = rx.adaptive_interpolate(bgr[i], bgr[i + 1], thresh, 0.5)
        assert abs(motion - gold[i]) <= 1e-4 and used == bool(gold[i] > thresh)
        gray = [cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in (bgr[i], bgr[i + 1])]
        assert abs(rx.calc_motion(cv2.resize(gray[0], (480, 270)), cv2.resize(gray[1], (480, 270))) - gold[i]) <= 1e-4
        if not used:                                                           # :866-868 linear blend, bit-exact
            for k, t in enumerate((0.25, 0.5, 0.75)):
                assert np.array_equal(frames[k], cv2.addWeighted(bgr[i], 1 - t, bgr[i + 1], t, 0))
        assert len(frames) == 3 and frames[0].shape == bgr[i].shape
    for i in range(2):
        gray = [cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in (bgr[i], bgr[i + 1])]
        c, flow = rx.estimate_motion_complexity(gray[0], gray[1])
        assert abs(c - g["complexity"][i]) <= 1e-5 and rx.mcar_tier(c) == "linear" and flow.shape == gray[0].shape + (2,)


def test_rife_interpolate_call_site(engine, golden, weights):
    """BGR in/out, cv2.BORDER_REFLECT padding to 128, scale 0.5 / 1.0: against the reference's own rife_interpolate lines run
    on the oracle IFNet.  bf16 kernels: <= 2e-2 of the range and PSNR >= 45 dB; fp32 kernels: <= 1 level."""
    g = golden("experiments.npz")
    n, H, W, seed = g["rife_clip"].tolist()
    small = np.ascontiguousarray(gen_clip(n, H, W, seed=seed, cut_every=0, fast=None)[..., ::-1])
    rx.RIFEModel._instance = engine
    for prec, tol in ((_abi.PRECISION_BF16, 6), (_abi.PRECISION_FP32, 1)):
        rx.RIFEModel.precision = prec
        for scale in (0.5, 1.0):
            for t in (0.25, 0.5):
                got = rx.rife_interpolate(small[0], small[1], t, scale)
                want = g[f"rife_s{scale}_t{t}"]
                d = np.abs(got.astype(int) - want.astype(int))
                assert d.max() <= tol, (prec, scale, t, d.max())
                if prec == _abi.PRECISION_BF16:
                    assert 10 * np.log10(255.0 ** 2 / max((d ** 2).mean(), 1e-12)) >= 45
    rx.RIFEModel.precision = _abi.PRECISION_BF16
    # flags really matter: RGB order or reflect-101 padding give a different frame
    a = engine.infer(small[0][None], small[1][None], [0.5], 0.5, _abi.PRECISION_FP32, flags=rx.FLAGS)[0, 0].cpu().numpy()
    b = engine.infer(small[0][None], small[1][None], [0.5], 0.5, _abi.PRECISION_FP32, flags=_abi.FLAG_BGR)[0, 0].cpu().numpy()
    c = engine.infer(np.ascontiguousarray(small[0][..., ::-1])[None], np.ascontiguousarray(small[1][..., ::-1])[None], [0.5], 0.5,
                     _abi.PRECISION_FP32, flags=_abi.FLAG_PAD_REFLECT_EDGE)[0, 0].cpu().numpy()
    assert not np.array_equal(a, b) and np.array_equal(a, c[..., ::-1])


@pytest.mark.parametrize("H,W", [(270, 480), (203, 317)])
def test_remap_flow_bit_exact_and_mcar_tiers(engine, H, W):
    """MCAR tier 2 (run_experiments.py:498-503): cv2.remap(prev, x + u t, y + v t, INTER_LINEAR) bit-exact with OpenCV's
    fixed-point path for a real Farneback field and for large / out-of-frame displacements; then the three tiers."""
    rgb = gen_clip(1, H, W, seed=6, cut_every=0, fast=(0, 2))
    bgr = np.ascontiguousarray(rgb[..., ::-1])
    g = [cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in bgr]
    flow = cv2.calcOpticalFlowFarneback(g[0], g[1], None, 0.5, 1, 15, 1, 5, 1.1, 0)
    rng = np.random.default_rng(0)
    wild = (rng.standard_normal(flow.shape) * 40).astype(np.float32)          # leaves the frame: BORDER_CONSTANT zeros
    y, x = np.mgrid[0:H, 0:W].astype(np.float32)
    for fl in (flow, wild):
        for t in (0.25, 0.5, 0.75):
            want = cv2.remap(bgr[0], x + fl[..., 0] * t, y + fl[..., 1] * t, cv2.INTER_LINEAR)
            assert np.array_equal(engine.remap_flow(bgr[0], fl, t).cpu().numpy(), want)
    gray1 = cv2.remap(g[0], x + flow[..., 0] * 0.5, y + flow[..., 1] * 0.5, cv2.INTER_LINEAR)
    assert np.array_equal(engine.remap_flow(g[0], flow, 0.5).cpu().numpy(), gray1)
    # tiers: thresholds placed around the pair's complexity
    c, dflow = rx.estimate_motion_complexity(g[0], g[1])
    lin, tier = rx.mcar_interpolate_cached(bgr[0], bgr[1], 0.5, c, dflow, 0.5, c + 0.1, c + 0.2)
    assert tier == "linear" and np.array_equal(lin, cv2.addWeighted(bgr[0], 0.5, bgr[1], 0.5, 0))
    fw, tier = rx.mcar_interpolate_cached(bgr[0], bgr[1], 0.25, c, dflow, 0.5, c - 0.01, c + 0.2)
    ref_flow = dflow.cpu().numpy()
    want = cv2.addWeighted(cv2.remap(bgr[0], x + ref_flow[..., 0] * 0.25, y + ref_flow[..., 1] * 0.25, cv2.INTER_LINEAR), 0.75, bgr[1], 0.25, 0)
    assert tier == "flow" and np.array_equal(fw, want)
    rx.RIFEModel._instance = engine
    rf, tier, c2 = rx.mcar_interpolate(bgr[0], bgr[1], 0.5, 1.0, -2.0, -1.0)
    assert tier == "rife" and rf.shape == bgr[0].shape and abs(c2 - c) < 1e-12
