"""Reduction kernels (PSNR, SSIM full/quarter-res, flow statistics, Laplacian variance, HUD mask, warp/blend)
against the oracle: integer results bit-exact, SSIM <= 1e-9, flow mean/std 1e-5 relative."""
import cv2
import numpy as np
import pytest
import torch

import vfi_b200  # noqa: F401
from oracle import ifnet_ref, metrics_ref, router_ref
from vfi_b200.engine import HudState
from vfi_b200.synth import gen_clip, noisy_copy

pytestmark = pytest.mark.gpu


def _sse(a, b):
    return int(((a.astype(np.int64) - b.astype(np.int64)) ** 2).sum())


@pytest.mark.parametrize("shape", [(3, 1080, 1920, 3), (5, 37, 53, 3), (2, 1, 1, 3), (4, 360, 640, 3)])
def test_psnr_exact(engine, shape):
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, shape, dtype=np.uint8)
    b = noisy_copy(a, seed=5)
    b[0] = a[0]                                                    # identical pair -> inf
    sse, ps = engine.psnr_sse(a, b)
    assert sse.tolist() == [_sse(a[i], b[i]) for i in range(shape[0])]
    ps = ps.cpu().numpy()
    assert ps[0] == np.inf
    for i in range(1, shape[0]):
        assert abs(ps[i] - metrics_ref.psnr_simple(a[i], b[i])) <= 1e-12 * abs(ps[i])


def test_psnr_extremes_and_unaligned(engine):
    a, b = np.zeros((1, 100, 100, 3), np.uint8), np.full((1, 100, 100, 3), 255, np.uint8)
    sse, ps = engine.psnr_sse(a, b)
    assert int(sse[0]) == 30000 * 65025 and float(ps[0]) == 0.0    # reference test_metrics.py:39-47
    big = torch.from_numpy(np.random.default_rng(2).integers(0, 256, (2, 1001), dtype=np.uint8)).cuda()
    ua, ub = big[:, 1:], big.flip(1)[:, 1:]                        # non-contiguous -> engine makes them contiguous
    sse, _ = engine.psnr_sse(ua, ub)
    assert sse.tolist() == [_sse(ua[i].cpu().numpy(), ub[i].cpu().numpy()) for i in range(2)]


@pytest.mark.parametrize("H,W", [(1080, 1920), (360, 640), (45, 71), (7, 7), (129, 250), (1440, 2560), (600, 129)])
def test_ssim_full_res(engine, H, W):
    fr = gen_clip(2, H, W, seed=3, cut_every=0, fast=None, hud=H > 100)
    pred = noisy_copy(fr, seed=6)
    got = engine.ssim(pred, fr, 1).cpu().numpy()
    want = np.array([metrics_ref.ssim_simple(pred[i], fr[i]) for i in range(len(fr))])
    assert np.abs(got - want).max() <= 1e-9
    assert abs(float(engine.ssim(fr[:1], fr[:1], 1)[0]) - 1.0) < 1e-12


@pytest.mark.parametrize("H,W", [(1080, 1920), (360, 640), (28, 28), (480, 640), (1000, 1777), (486, 482), (30, 31)])
def test_scene_cut_ssim_quarter_res(engine, H, W):
    fr = gen_clip(7, H, W, seed=3, cut_every=4, fast=None, hud=H > 100)
    got = engine.ssim(fr[:-1], fr[1:], 4).cpu().numpy()
    want = np.array([router_ref.scene_ssim(fr[i], fr[i + 1]) for i in range(len(fr) - 1)])
    assert np.abs(got - want).max() <= 1e-9
    assert np.array_equal(got < 0.65, want < 0.65)
    if H >= 360:
        assert (want < 0.65).any() and (want >= 0.65).any()


@pytest.mark.parametrize("shape", [(3, 1080, 1920, 3), (2, 360, 640, 3), (4, 45, 71, 3), (1, 7, 7, 3), (2, 130, 123, 3), (1, 2160, 3840, 3)])
def test_fused_psnr_ssim_equals_separate_kernels(engine, shape):
    """vfi_psnr_ssim_u8 (one read of the frames): SSE identical to vfi_psnr_u8 and to numpy, SSIM <= 1e-9 from the oracle."""
    n, H, W, _ = shape
    fr = gen_clip(n - 1, H, W, seed=13, cut_every=0, fast=None, hud=H > 100) if H >= 16 else \
        np.random.default_rng(1).integers(0, 256, shape, dtype=np.uint8)
    pred = noisy_copy(fr, seed=7, amp=5)
    pred[0] = fr[0]
    sse, ps, ss = engine.psnr_ssim(pred, fr)
    assert sse.tolist() == [_sse(pred[i], fr[i]) for i in range(n)] == engine.psnr_sse(pred, fr)[0].tolist()
    assert float(ps[0]) == np.inf and abs(float(ss[0]) - 1.0) < 1e-12
    want = np.array([metrics_ref.ssim_simple(pred[i], fr[i]) for i in range(n)])
    assert np.abs(ss.cpu().numpy() - want).max() <= 1e-9
    for i in range(1, n):
        assert abs(float(ps[i]) - metrics_ref.psnr_simple(pred[i], fr[i])) <= 1e-12 * abs(float(ps[i]))


def test_ssim_random_noise_range(engine):                          # reference test_metrics.py:78-85
    rng = np.random.default_rng(4)
    a, b = (rng.integers(0, 256, (1, 100, 100, 3), dtype=np.uint8) for _ in range(2))
    s = float(engine.ssim(a, b, 1)[0])
    assert -1 <= s <= 1 and abs(s - metrics_ref.ssim_simple(a[0], b[0])) < 1e-9


@pytest.mark.parametrize("H,W", [(360, 640), (1080, 1920)])
def test_flow_stats(engine, H, W):
    fr = gen_clip(2, H, W, seed=8, cut_every=0, fast=None)
    flows = np.stack([router_ref.farneback(cv2.cvtColor(fr[i], cv2.COLOR_RGB2GRAY), cv2.cvtColor(fr[i + 1], cv2.COLOR_RGB2GRAY))
                      for i in range(2)])
    got = engine.flow_stats(flows).cpu().numpy()
    for i in range(2):
        m, mx, sd, _ = router_ref.flow_stats(flows[i])
        assert abs(got[i, 0] - m) <= 1e-5 * abs(m) and got[i, 1] == mx and abs(got[i, 2] - sd) <= 1e-5 * abs(sd)


def test_flow_stats_synthetic_extremes(engine):
    f = np.zeros((1, 64, 96, 2), np.float32)
    assert engine.flow_stats(f).cpu().numpy().tolist() == [[0.0, 0.0, 0.0]]
    f[0, 10, 20] = (30.0, 40.0)
    g = engine.flow_stats(f).cpu().numpy()[0]
    assert g[1] == 50.0 and abs(g[0] - 50.0 / (64 * 96)) < 1e-12


@pytest.mark.parametrize("H,W", [(1080, 1920), (360, 640), (33, 65), (2, 2), (2160, 3840), (700, 121), (5, 300)])
def test_laplacian_var(engine, H, W):
    rng = np.random.default_rng(9)
    fr = rng.integers(0, 256, (3, H, W, 3), dtype=np.uint8)
    got = engine.laplacian_var(fr).cpu().numpy()
    want = np.array([router_ref.laplacian_var(f) for f in fr])
    assert np.allclose(got, want, rtol=1e-12, atol=1e-9)


@pytest.mark.parametrize("H,W", [(360, 640), (1080, 1920), (1440, 2560), (480, 640), (720, 1280), (1000, 1777), (540, 960),
                                 (480, 854), (100, 100)])
def test_hud_mask_bit_exact(engine, H, W):
    """Any frame size the reference's cv2.resize calls accept (adaptive_pipeline.py:203-204), incl. the reference's own
    480x640 fixture (tests/conftest.py:15-18), non-integer ratios and frames smaller than the 320x180 thumbnail."""
    clip = gen_clip(11 if H <= 360 else 6, H, W, seed=4, cut_every=0, fast=None)
    ref, hs = router_ref.RouterRef(), HudState(engine)
    seen_mask = False
    for f in clip:
        m_ref, c_ref = ref.detect_hud(f)
        m, c = hs.update(f)
        assert np.array_equal(m.cpu().numpy().astype(bool), m_ref) and float(c[0]) == c_ref
        seen_mask |= bool(m_ref.any())
    assert seen_mask and len(hs) == min(len(clip), 10)


def test_hud_thumbnail_ring_matches_cv2(engine):
    """The ring holds cv2.resize(gray, (320, 180)) bit for bit at a non-integer ratio (general INTER_LINEAR taps)."""
    fr = np.random.default_rng(3).integers(0, 256, (3, 500, 888, 3), dtype=np.uint8)
    hs = HudState(engine)
    for f in fr:
        hs.update(f)
    ring = hs.ring.cpu().numpy()
    for i, f in enumerate(fr):
        assert np.array_equal(ring[i], cv2.resize(cv2.cvtColor(f, cv2.COLOR_RGB2GRAY), (320, 180)))


@pytest.mark.parametrize("H,W", [(128, 192), (1080, 1920), (33, 47)])
def test_warp_blend_vs_oracle(engine, H, W):
    """north_star kernel (b) on its own: grid_sample(border, align_corners=True) x2 + sigmoid blend + u8 store."""
    rng = np.random.default_rng(11)
    fr = gen_clip(1, H, W, seed=12, cut_every=0, fast=None, hud=False)
    flow = (rng.standard_normal((1, H, W, 4)) * 6).astype(np.float32)
    flow[0, :4] *= 20                                              # far out of frame: border clamp
    mask = (rng.standard_normal((1, H, W)) * 3).astype(np.float32)
    out, of = engine.warp_blend(fr[:1], fr[1:2], torch.from_numpy(flow), torch.from_numpy(mask), want_float=True)
    i0, i1 = ifnet_ref.to_tensor(fr[0]), ifnet_ref.to_tensor(fr[1])
    fl = torch.from_numpy(flow).permute(0, 3, 1, 2)
    m = torch.sigmoid(torch.from_numpy(mask)[:, None])
    want = ifnet_ref.warp(i0, fl[:, :2]) * m + ifnet_ref.warp(i1, fl[:, 2:4]) * (1 - m)
    want = want[0].permute(1, 2, 0).numpy()
    assert float(np.abs(of[0].cpu().numpy() - want).max()) <= 2e-5
    assert np.abs(out[0].cpu().numpy().astype(int) - (want * 255).clip(0, 255).astype(np.uint8).astype(int)).max() <= 1


@pytest.mark.parametrize("H,W", [(1080, 1920), (360, 640), (6, 10), (270, 482), (2, 2)])
def test_yuv420_bit_exact_vs_cv2(engine, H, W):
    """vfi_rgb_to_yuv420_u8 = cv2.cvtColor(COLOR_RGB2YUV_I420 / COLOR_BGR2YUV_I420), bit for bit (aligned fast path: W % 4 == 0;
    byte path otherwise)."""
    rgb = np.random.default_rng(H + W).integers(0, 256, (3, H, W, 3), dtype=np.uint8)
    got = engine.to_yuv420(rgb).cpu().numpy()
    assert got.shape == (3, H * 3 // 2, W)
    for i in range(3):
        assert np.array_equal(got[i], cv2.cvtColor(rgb[i], cv2.COLOR_RGB2YUV_I420))
    assert np.array_equal(engine.to_yuv420(rgb[0], bgr=True).cpu().numpy(), cv2.cvtColor(rgb[0], cv2.COLOR_BGR2YUV_I420))
    ext = np.zeros((2, H, W, 3), np.uint8)
    ext[1] = 255                                                        # black / white: Y 16 / 235, chroma 128
    e = engine.to_yuv420(ext).cpu().numpy()
    assert e[0, :H].max() == 16 and e[1, :H].min() == 235 and (e[:, H:] == 128).all()


def test_yuv420_rejects_odd_sizes(engine):
    from vfi_b200 import _abi
    with pytest.raises(_abi.VFIError, match="even frame sizes"):
        engine.to_yuv420(np.zeros((1, 5, 8, 3), np.uint8))
