"""CPU oracle: PSNR / SSIM of evaluation/metrics.py, restated on numpy + scipy + cv2.

TEST INFRASTRUCTURE ONLY (see oracle/ifnet_ref.py header for who may import it).

Pinned by: the reference's analytic known-answer tests (tests/test_metrics.py:21-25
identical -> inf, :39-47 0 vs 255 -> 0 dB, :72-76 identical -> SSIM 1, :49-57/:98-106
symmetry) which tests/test_oracle_metrics.py replays, and for PSNR by importing the
reference's own compute_psnr_simple in this container (tests/golden/make_golden.py).
SSIM: scikit-image >= 0.21 (requirements.txt:49) is absent here, so
structural_similarity is restated from its published algorithm;
PARITY UNPINNED against skimage itself.
"""
import cv2
import numpy as np
from scipy.ndimage import uniform_filter


def rgb2gray_u8(rgb):
    """cv2.cvtColor(RGB2GRAY) integer identity, checked against cv2 in tests."""
    r = rgb[..., 0].astype(np.int64)
    g = rgb[..., 1].astype(np.int64)
    b = rgb[..., 2].astype(np.int64)
    return ((r * 9798 + g * 19235 + b * 3735 + 16384) >> 15).astype(np.uint8)


def psnr_simple(pred, gt):
    """evaluation/metrics.py:299-315."""
    if pred.shape != gt.shape:
        pred = cv2.resize(pred, (gt.shape[1], gt.shape[0]))
    mse = np.mean((pred.astype(float) - gt.astype(float)) ** 2)
    if mse == 0:
        return float("inf")
    return float(10 * np.log10(255.0 ** 2 / mse))


def ssim_gray_u8(a, b):
    """skimage.metrics.structural_similarity(a, b) for 2-D uint8, all defaults (App. B.2)."""
    assert a.ndim == 2 and a.shape == b.shape and a.dtype == np.uint8
    win = 7
    cov_norm = win * win / (win * win - 1.0)
    x = a.astype(np.float64)
    y = b.astype(np.float64)
    ux = uniform_filter(x, size=win)
    uy = uniform_filter(y, size=win)
    uxx = uniform_filter(x * x, size=win)
    uyy = uniform_filter(y * y, size=win)
    uxy = uniform_filter(x * y, size=win)
    vx = cov_norm * (uxx - ux * ux)
    vy = cov_norm * (uyy - uy * uy)
    vxy = cov_norm * (uxy - ux * uy)
    c1 = (0.01 * 255) ** 2
    c2 = (0.03 * 255) ** 2
    s = ((2 * ux * uy + c1) * (2 * vxy + c2)) / ((ux * ux + uy * uy + c1) * (vx + vy + c2))
    pad = (win - 1) // 2
    return float(s[pad:-pad, pad:-pad].mean(dtype=np.float64))


def ssim_gray_u8_exact(a, b):
    """Same quantity from exact integer 7x7 window sums (the formulation the CUDA kernel uses).

    Agreement with ssim_gray_u8 (uniform_filter route) to ~1e-12 is asserted in the tests;
    it shows the integer restatement is the same function, not a different SSIM.
    """
    a = a.astype(np.int64)
    b = b.astype(np.int64)

    def box(v):
        c = np.cumsum(np.cumsum(np.pad(v, ((1, 0), (1, 0))), 0), 1)
        return c[7:, 7:] - c[:-7, 7:] - c[7:, :-7] + c[:-7, :-7]

    sa, sb, saa, sbb, sab = box(a), box(b), box(a * a), box(b * b), box(a * b)
    n = 49.0
    ux = sa / n
    uy = sb / n
    cov_norm = 49.0 / 48.0
    vx = cov_norm * (saa / n - ux * ux)
    vy = cov_norm * (sbb / n - uy * uy)
    vxy = cov_norm * (sab / n - ux * uy)
    c1 = (0.01 * 255) ** 2
    c2 = (0.03 * 255) ** 2
    s = ((2 * ux * uy + c1) * (2 * vxy + c2)) / ((ux * ux + uy * uy + c1) * (vx + vy + c2))
    return float(s.mean(dtype=np.float64))


def ssim_simple(pred, gt):
    """evaluation/metrics.py:318-335."""
    if pred.shape != gt.shape:
        pred = cv2.resize(pred, (gt.shape[1], gt.shape[0]))
    if pred.ndim == 3:
        pg = cv2.cvtColor(pred, cv2.COLOR_RGB2GRAY)
        gg = cv2.cvtColor(gt, cv2.COLOR_RGB2GRAY)
    else:
        pg, gg = pred, gt
    return ssim_gray_u8(pg, gg)
