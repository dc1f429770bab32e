"""CPU oracle: cv2.resize(uint8, INTER_LINEAR) restated in integer arithmetic.

TEST INFRASTRUCTURE ONLY (see oracle/ifnet_ref.py header for who may import it).

The router calls cv2.resize twice per frame (reference models/novel/adaptive_pipeline.py:149-155 `fx=fy=0.25`,
:203-204 `(320, 180)`); OpenCV is third-party (opencv-python-headless, requirements.txt:15), so its published
algorithm is restated here (imgproc/resize.cpp, resizeGeneric_ with HResizeLinear<uchar,int,short,2048> and
VResizeLinear<uchar,int,short,FixedPtCast<int,uchar,22>>) and PINNED against the cv2 4.13.0 in this image by
tests/test_oracle_router.py::test_resize_restatement_matches_cv2 (bit-exact on down- and up-scaling, odd sizes).
The CUDA kernels (csrc/reduce.cu lin_tap / resized_gray) implement exactly this function.
"""
import numpy as np

COEF_BITS = 11
ONE = 1 << COEF_BITS


def _taps(dst, src, inv_scale, clamp_frac):
    """Per destination index: (s0, s1, a0, a1).  resize.cpp computes fx in float from a double product."""
    scale = 1.0 / inv_scale
    d = np.arange(dst, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int64)
    f = (f - s.astype(np.float32)).astype(np.float32)
    if clamp_frac:                       # x axis: `if (sx < 0) fx = 0, sx = 0; if (sx >= ssize.width-1) fx = 0, sx = width-1`
        lo, hi = s < 0, s >= src - 1
        f = np.where(lo | hi, np.float32(0), f)
        s = np.where(lo, 0, np.where(hi, src - 1, s))
    a0 = np.rint((np.float32(1) - f) * np.float32(ONE)).astype(np.int64)
    a1 = np.rint(f * np.float32(ONE)).astype(np.int64)
    return np.clip(s, 0, src - 1), np.clip(s + 1, 0, src - 1), a0, a1


def resize_linear_u8(gray, dsize=None, fx=None, fy=None):
    """cv2.resize(gray, dsize) or cv2.resize(gray, None, fx=fx, fy=fy) for a 2-D uint8 image, INTER_LINEAR."""
    H, W = gray.shape
    if dsize is not None:
        dw, dh = dsize
        isx, isy = dw / W, dh / H
    else:
        dw, dh = int(np.rint(W * fx)), int(np.rint(H * fy))     # saturate_cast<int> = round half to even
        isx, isy = fx, fy
    x0, x1, ax0, ax1 = _taps(dw, W, isx, True)
    y0, y1, ay0, ay1 = _taps(dh, H, isy, False)                 # y: rows are clipped, the fraction is kept
    g = gray.astype(np.int64)
    hr = g[:, x0] * ax0[None] + g[:, x1] * ax1[None]
    s0, s1 = hr[y0], hr[y1]
    out = (((ay0[:, None] * (s0 >> 4)) >> 16) + ((ay1[:, None] * (s1 >> 4)) >> 16) + 2) >> 2
    return out.astype(np.uint8)
