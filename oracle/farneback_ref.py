"""CPU oracle: OpenCV's dense Farneback optical flow restated in numpy.

TEST INFRASTRUCTURE ONLY (see oracle/ifnet_ref.py header for who may import it).

The reference calls `cv2.calcOpticalFlowFarneback(g0, g1, None, pyr_scale=0.5, levels=3, winsize=15, iterations=3,
poly_n=5, poly_sigma=1.2, flags=0)` (models/novel/adaptive_pipeline.py:119-128; same parameters at
evaluation/metrics.py:379-381).  OpenCV is a third-party dependency that is not vendored in /root/reference
(opencv-python-headless >= 4.8, requirements.txt:15; 4.13.0 in this image), so its published algorithm
(modules/video/src/optflowgf.cpp: calcOpticalFlowFarneback -> FarnebackPolyExp / FarnebackUpdateMatrices /
FarnebackUpdateFlow_Blur, with imgproc's GaussianBlur and resize) is restated here, operation by operation and
in the same floating-point types, and PINNED against cv2 itself by tests/test_oracle_farneback.py (max |diff|
<= 1e-3 px; typically 1e-5).  The CUDA path (csrc/farneback.cu) is compared with this restatement and with
cv2-generated golden flow statistics.

flags = 0 only (box window, zero initial flow): that is the only mode the reference uses.
"""
import numpy as np

F = np.float32
MIN_SIZE = 32


def cv_round(x):
    """cvRound: round half to even."""
    return int(np.rint(x))


def gaussian_kernel(n, sigma):
    """cv::getGaussianKernel(n, sigma, CV_32F) (imgproc/smooth.dispatch.cpp)."""
    small = {1: [1.0], 3: [0.25, 0.5, 0.25], 5: [0.0625, 0.25, 0.375, 0.25, 0.0625],
             7: [0.03125, 0.109375, 0.21875, 0.28125, 0.21875, 0.109375, 0.03125]}
    if sigma <= 0 and n in small:
        return np.array(small[n], F)
    s = sigma if sigma > 0 else ((n - 1) * 0.5 - 1) * 0.3 + 0.8
    x = np.arange(n, dtype=np.float64) - (n - 1) * 0.5
    k = np.exp(-0.5 / (s * s) * x * x)
    return (k / k.sum()).astype(F)


def _reflect101(i, n):
    i = np.abs(i)
    i = np.where(i >= n, 2 * (n - 1) - i, i)
    return np.clip(i, 0, n - 1)


def gaussian_blur(img, ksize, sigma):
    """cv::GaussianBlur on CV_32F, BORDER_REFLECT_101: separable, symmetric pairs summed first."""
    k = gaussian_kernel(ksize, sigma)
    r = ksize // 2
    H, W = img.shape
    xi = np.arange(W)
    row = img * k[r]
    for d in range(1, r + 1):
        row = row + k[r + d] * (img[:, _reflect101(xi - d, W)] + img[:, _reflect101(xi + d, W)])
    yi = np.arange(H)
    out = row * k[r]
    for d in range(1, r + 1):
        out = out + k[r + d] * (row[_reflect101(yi - d, H)] + row[_reflect101(yi + d, H)])
    return out.astype(F)


def _lin_taps(dst, src, inv_scale, clamp_frac):
    scale = 1.0 / inv_scale
    f = ((np.arange(dst, dtype=np.float64) + 0.5) * scale - 0.5).astype(F)
    s = np.floor(f).astype(np.int64)
    f = (f - s.astype(F)).astype(F)
    if clamp_frac:
        lo, hi = s < 0, s >= src - 1
        f = np.where(lo | hi, F(0), f)
        s = np.where(lo, 0, np.where(hi, src - 1, s))
    return np.clip(s, 0, src - 1), np.clip(s + 1, 0, src - 1), (F(1) - f).astype(F), f


def resize_linear_f32(img, dw, dh):
    """cv::resize(CV_32F[Cn], Size(dw, dh), INTER_LINEAR) (imgproc/resize.cpp); img: (H, W) or (H, W, C)."""
    H, W = img.shape[:2]
    if (dw, dh) == (W, H):
        return img.copy()
    x0, x1, a0, a1 = _lin_taps(dw, W, dw / W, True)
    y0, y1, b0, b1 = _lin_taps(dh, H, dh / H, False)
    ex = (slice(None),) + (None,) * (img.ndim - 2)
    hr = img[:, x0] * a0[ex] + img[:, x1] * a1[ex]
    ey = (slice(None), None) + (None,) * (img.ndim - 2)
    return (hr[y0] * b0[ey] + hr[y1] * b1[ey]).astype(F)


def prepare_gaussian(n, sigma):
    """FarnebackPrepareGaussian."""
    if sigma < np.finfo(F).eps:
        sigma = n * 0.3
    x = np.arange(-n, n + 1)
    g = np.exp(-x * x / (2 * sigma * sigma)).astype(F)
    s = 1.0 / g.astype(np.float64).sum()            # double accumulator over the float taps
    g = (g.astype(np.float64) * s).astype(F)
    xg = (x * g.astype(np.float64)).astype(F)
    xxg = (x * x * g.astype(np.float64)).astype(F)
    G = np.zeros((6, 6))
    gd = g.astype(np.float64)
    for yy in range(-n, n + 1):
        for xx in range(-n, n + 1):
            w = F(g[yy + n] * g[xx + n])             # float product, accumulated in double
            G[0, 0] += w
            G[1, 1] += w * xx * xx
            G[3, 3] += w * xx * xx * xx * xx
            G[5, 5] += w * xx * xx * yy * yy
    G[2, 2] = G[0, 3] = G[0, 4] = G[3, 0] = G[4, 0] = G[1, 1]
    G[4, 4] = G[3, 3]
    G[3, 4] = G[4, 3] = G[5, 5]
    inv = np.linalg.inv(G)
    return g[n:], xg[n:], xxg[n:], inv[1, 1], inv[0, 3], inv[3, 3], inv[5, 5]


def poly_exp(src, n=5, sigma=1.2):
    """FarnebackPolyExp: (H, W) f32 -> (H, W, 5) f32 polynomial coefficients."""
    g, xg, xxg, ig11, ig03, ig33, ig55 = prepare_gaussian(n, sigma)
    H, W = src.shape
    yi = np.arange(H)
    r0 = src * g[0]
    r1 = np.zeros_like(src)
    r2 = np.zeros_like(src)
    for k in range(1, n + 1):
        s0 = src[np.maximum(yi - k, 0)]
        s1 = src[np.minimum(yi + k, H - 1)]
        p = s0 + s1
        r0 = r0 + g[k] * p
        r1 = r1 + xg[k] * (s1 - s0)
        r2 = r2 + xxg[k] * p
    xi = np.arange(W)

    def at(r, d):
        return r[:, np.clip(xi + d, 0, W - 1)]

    D = np.float64
    b1 = (r0 * g[0]).astype(D)
    b3 = (r1 * g[0]).astype(D)
    b5 = (r2 * g[0]).astype(D)
    b2 = np.zeros((H, W), D)
    b4 = np.zeros((H, W), D)
    b6 = np.zeros((H, W), D)
    for k in range(1, n + 1):
        p0, m0 = at(r0, k), at(r0, -k)
        p1, m1 = at(r1, k), at(r1, -k)
        p2, m2 = at(r2, k), at(r2, -k)
        tg = (p0 + m0).astype(D)                      # float sum widened
        b1 += tg * D(g[k])
        b4 += tg * D(xxg[k])
        b2 += ((p0 - m0) * xg[k]).astype(D)           # float products widened
        b3 += ((p1 + m1) * g[k]).astype(D)
        b6 += ((p1 - m1) * xg[k]).astype(D)
        b5 += ((p2 + m2) * g[k]).astype(D)
    out = np.empty((H, W, 5), F)
    out[..., 1] = (b2 * ig11).astype(F)
    out[..., 0] = (b3 * ig11).astype(F)
    out[..., 3] = (b1 * ig03 + b4 * ig33).astype(F)
    out[..., 2] = (b1 * ig03 + b5 * ig33).astype(F)
    out[..., 4] = (b6 * ig55).astype(F)
    return out


BORDER = np.array([0.14, 0.14, 0.4472, 0.4472, 0.4472], F)


def update_matrices(R0, R1, flow):
    """FarnebackUpdateMatrices: (H, W, 5) x2 + (H, W, 2) -> M (H, W, 5), all float32."""
    H, W = flow.shape[:2]
    xg, yg = np.meshgrid(np.arange(W), np.arange(H))
    dx, dy = flow[..., 0], flow[..., 1]
    fx = xg.astype(F) + dx
    fy = yg.astype(F) + dy
    x1 = np.floor(fx).astype(np.int64)
    y1 = np.floor(fy).astype(np.int64)
    fx = fx - x1.astype(F)
    fy = fy - y1.astype(F)
    inside = (x1 >= 0) & (x1 < W - 1) & (y1 >= 0) & (y1 < H - 1)
    xc, yc = np.clip(x1, 0, W - 2), np.clip(y1, 0, H - 2)
    a00 = (F(1) - fx) * (F(1) - fy)
    a01 = fx * (F(1) - fy)
    a10 = (F(1) - fx) * fy
    a11 = fx * fy
    p00, p01, p10, p11 = R1[yc, xc], R1[yc, xc + 1], R1[yc + 1, xc], R1[yc + 1, xc + 1]
    r = a00[..., None] * p00 + a01[..., None] * p01 + a10[..., None] * p10 + a11[..., None] * p11
    r2 = np.where(inside, r[..., 0], F(0))
    r3 = np.where(inside, r[..., 1], F(0))
    r4 = np.where(inside, (R0[..., 2] + r[..., 2]) * F(0.5), R0[..., 2])
    r5 = np.where(inside, (R0[..., 3] + r[..., 3]) * F(0.5), R0[..., 3])
    r6 = np.where(inside, (R0[..., 4] + r[..., 4]) * F(0.25), R0[..., 4] * F(0.5))
    r2 = (R0[..., 0] - r2) * F(0.5)
    r3 = (R0[..., 1] - r3) * F(0.5)
    r2 = r2 + r4 * dy + r6 * dx
    r3 = r3 + r6 * dy + r5 * dx

    def edge(n):
        idx = np.arange(n)
        lo = np.where(idx < 5, BORDER[np.clip(idx, 0, 4)], F(1)).astype(F)
        hi = np.where(idx >= n - 5, BORDER[np.clip(n - idx - 1, 0, 4)], F(1)).astype(F)
        return lo, hi
    sx0, sx1 = edge(W)
    sy0, sy1 = edge(H)
    scale = ((sx0 * sx1)[None, :] * sy0[:, None]) * sy1[:, None]
    r2, r3, r4, r5, r6 = r2 * scale, r3 * scale, r4 * scale, r5 * scale, r6 * scale
    M = np.empty((H, W, 5), F)
    M[..., 0] = r4 * r4 + r6 * r6
    M[..., 1] = (r4 + r5) * r6
    M[..., 2] = r5 * r5 + r6 * r6
    M[..., 3] = r4 * r2 + r6 * r3
    M[..., 4] = r6 * r2 + r5 * r3
    return M


def update_flow_blur(M, block_size=15):
    """The blur + 2x2 solve of FarnebackUpdateFlow_Blur: clamped box window sums in double -> flow (H, W, 2) f32.
    (OpenCV keeps running sums; a direct cumulative sum differs from them only in the last double bits.)"""
    H, W = M.shape[:2]
    m = block_size // 2
    D = np.float64
    Md = M.astype(D)
    yi = np.arange(H)
    cs = np.concatenate([np.zeros((1, W, 5), D), np.cumsum(Md, 0)], 0)
    v = np.zeros((H, W, 5), D)
    # clamped window [y-m, y+m]: rows below 0 count as row 0, rows above H-1 as row H-1
    lo, hi = yi - m, yi + m
    v = cs[np.clip(hi, 0, H - 1) + 1] - cs[np.clip(lo, 0, H - 1)]
    v += np.maximum(0, -lo)[:, None, None] * Md[0][None] + np.maximum(0, hi - (H - 1))[:, None, None] * Md[H - 1][None]
    xi = np.arange(W)
    cs = np.concatenate([np.zeros((H, 1, 5), D), np.cumsum(v, 1)], 1)
    lo, hi = xi - m, xi + m
    g = cs[:, np.clip(hi, 0, W - 1) + 1] - cs[:, np.clip(lo, 0, W - 1)]
    g += np.maximum(0, -lo)[None, :, None] * v[:, :1] + np.maximum(0, hi - (W - 1))[None, :, None] * v[:, W - 1:]
    g *= 1.0 / (block_size * block_size)
    g11, g12, g22, h1, h2 = (g[..., i] for i in range(5))
    idet = 1.0 / (g11 * g22 - g12 * g12 + 1e-3)
    flow = np.empty((H, W, 2), F)
    flow[..., 0] = ((g11 * h2 - g12 * h1) * idet).astype(F)
    flow[..., 1] = ((g22 * h1 - g12 * h2) * idet).astype(F)
    return flow


def pyramid_levels(W, H, levels=3, pyr_scale=0.5):
    k, scale = 0, 1.0
    while k < levels:
        scale *= pyr_scale
        if W * scale < MIN_SIZE or H * scale < MIN_SIZE:
            break
        k += 1
    return k


def level_geometry(W, H, k, pyr_scale=0.5):
    """(scale, sigma, smooth_sz, width, height) of pyramid level k (0 = full resolution)."""
    scale = 1.0
    for _ in range(k):
        scale *= pyr_scale
    sigma = (1.0 / scale - 1) * 0.5
    smooth = max(cv_round(sigma * 5) | 1, 3)
    return scale, sigma, smooth, cv_round(W * scale), cv_round(H * scale)


def farneback(gray0, gray1, pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2):
    """calcOpticalFlowFarneback(..., flags=0) for two (H, W) uint8 images -> (H, W, 2) float32."""
    H, W = gray0.shape
    imgs = [gray0.astype(F), gray1.astype(F)]
    flow = None
    for k in range(pyramid_levels(W, H, levels, pyr_scale), -1, -1):
        _, sigma, smooth, w, h = level_geometry(W, H, k, pyr_scale)
        if flow is None:
            flow = np.zeros((h, w, 2), F)
        else:
            flow = (resize_linear_f32(flow, w, h) * F(1.0 / pyr_scale)).astype(F)
        R = [poly_exp(resize_linear_f32(gaussian_blur(im, smooth, sigma), w, h), poly_n, poly_sigma) for im in imgs]
        M = update_matrices(R[0], R[1], flow)
        for i in range(iterations):
            flow = update_flow_blur(M, winsize)
            if i < iterations - 1:
                M = update_matrices(R[0], R[1], flow)
    return flow
