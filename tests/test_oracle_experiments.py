"""oracle/experiments_ref.py (the restated experiment-runner call sites, rows N2 / N3) pinned against vectors produced by
the reference's own scripts/run_experiments.py functions (tests/golden/make_golden_experiments.py)."""
import cv2
import numpy as np

import vfi_b200  # noqa: F401
from oracle import experiments_ref as xr
from vfi_b200.synth import gen_clip, noisy_copy


def _clip(g):
    n, H, W, seed = g["clip"].tolist()
    rgb = gen_clip(n, H, W, seed=seed, cut_every=0, fast=tuple(g["fast"].tolist()))
    return np.ascontiguousarray(rgb[..., ::-1]), np.ascontiguousarray(noisy_copy(rgb, seed=5, amp=4)[..., ::-1])


def test_global_metrics_and_motion_match_reference(golden):
    g = golden("experiments.npz")
    bgr, pred = _clip(g)
    n = len(bgr) - 1
    ps = [xr.calc_psnr(pred[i], bgr[i]) for i in range(n + 1)] + [xr.calc_psnr(bgr[i], bgr[i + 1]) for i in range(n)]
    ss = [xr.calc_ssim(pred[i], bgr[i]) for i in range(n + 1)] + [xr.calc_ssim(bgr[i], bgr[i + 1]) for i in range(n)]
    assert ps == g["psnr"].tolist() and ss == g["ssim"].tolist()            # same expressions: bit-identical
    assert xr.calc_psnr(bgr[0], bgr[0]) == float("inf") and xr.calc_ssim(bgr[0], bgr[0]) == 1.0
    mo = [xr.calc_motion(*xr.thumbnails(bgr[i], bgr[i + 1])) for i in range(n)]
    assert mo == g["motion"].tolist()
    gray = [cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in bgr]
    cxs = [float(xr.estimate_motion_complexity(gray[i], gray[i + 1])[0]) for i in range(2)]
    assert cxs == g["complexity"][:2].tolist()


def test_rife_interpolate_restated_call_site(golden, oracle_model):
    g = golden("experiments.npz")
    n, H, W, seed = g["rife_clip"].tolist()
    small = np.ascontiguousarray(gen_clip(n, H, W, seed=seed, cut_every=0, fast=None)[..., ::-1])
    for scale in (0.5, 1.0):
        for t in (0.25, 0.5):
            got = xr.rife_interpolate(oracle_model, small[0], small[1], t, scale)
            assert np.array_equal(got, g[f"rife_s{scale}_t{t}"])
