"""Oracle PSNR/SSIM (oracle/metrics_ref.py) pinned against the reference's own known-answer tests
(/root/reference tests/test_metrics.py, replayed with seeds) and the reference-generated golden PSNRs."""
import cv2
import numpy as np
import pytest

import vfi_b200  # noqa: F401
from oracle import metrics_ref
from vfi_b200.synth import gen_clip, noisy_copy


def _rand(shape, seed):
    return np.random.default_rng(seed).integers(0, 256, shape, dtype=np.uint8)


class TestPSNR:  # reference tests/test_metrics.py:18-66
    def test_identical_inf(self):
        a = _rand((100, 100, 3), 0)
        assert metrics_ref.psnr_simple(a, a) == float("inf")

    def test_noise_range(self):
        a = _rand((100, 100, 3), 1)
        n = np.random.default_rng(2).integers(-10, 11, a.shape, dtype=np.int16)
        b = np.clip(a.astype(np.int16) + n, 0, 255).astype(np.uint8)
        assert 20 < metrics_ref.psnr_simple(a, b) < 50

    def test_black_white_is_0db(self):
        a, b = np.zeros((100, 100, 3), np.uint8), np.full((100, 100, 3), 255, np.uint8)
        assert metrics_ref.psnr_simple(a, b) == 0.0

    def test_symmetry(self):
        a, b = _rand((100, 100, 3), 3), _rand((100, 100, 3), 4)
        assert abs(metrics_ref.psnr_simple(a, b) - metrics_ref.psnr_simple(b, a)) < 1e-3

    def test_resize(self):
        a, b = _rand((50, 50, 3), 5), _rand((100, 100, 3), 6)
        assert np.isfinite(metrics_ref.psnr_simple(a, b))

    def test_golden_from_reference(self, golden):
        g = golden("psnr.npz")
        n, H, W, seed = g["clip"].tolist()
        fr = gen_clip(n, H, W, seed=seed, cut_every=0, fast=None)
        pred = noisy_copy(fr, seed=5, amp=3)
        got = [metrics_ref.psnr_simple(pred[i], fr[i]) for i in range(len(fr))]
        assert got == g["psnr_noise"].tolist()                                  # bit-identical float64
        got = [metrics_ref.psnr_simple(fr[i], fr[i + 1]) for i in range(len(fr) - 1)]
        assert got == g["psnr_cross"].tolist()
        assert metrics_ref.psnr_simple(fr[0][::2, ::2].copy(), fr[0]) == float(g["psnr_resize"][0])


class TestSSIM:  # reference tests/test_metrics.py:69-106
    def test_identical_is_one(self):
        a = _rand((100, 100, 3), 0)
        assert abs(metrics_ref.ssim_simple(a, a) - 1.0) < 1e-3

    def test_range(self):
        a, b = _rand((100, 100, 3), 1), _rand((100, 100, 3), 2)
        assert -1 <= metrics_ref.ssim_simple(a, b) <= 1

    def test_similar_high(self):
        a = np.random.default_rng(3).integers(100, 150, (100, 100, 3)).astype(np.uint8)
        n = np.random.default_rng(4).integers(-5, 6, a.shape, dtype=np.int16)
        b = np.clip(a.astype(np.int16) + n, 0, 255).astype(np.uint8)
        assert metrics_ref.ssim_simple(a, b) > 0.9

    def test_symmetry(self):
        a, b = _rand((100, 100, 3), 5), _rand((100, 100, 3), 6)
        assert abs(metrics_ref.ssim_simple(a, b) - metrics_ref.ssim_simple(b, a)) < 1e-3

    @pytest.mark.parametrize("shape", [(7, 7), (9, 23), (64, 48), (180, 320)])
    def test_integer_window_restatement_is_same_function(self, shape):
        a, b = _rand(shape, 7), _rand(shape, 8)
        assert abs(metrics_ref.ssim_gray_u8(a, b) - metrics_ref.ssim_gray_u8_exact(a, b)) < 1e-11

    def test_cross_metric_sanity(self):  # reference :109-142
        a = _rand((100, 100, 3), 9)
        n = np.random.default_rng(10).integers(-3, 4, a.shape, dtype=np.int16)
        small = np.clip(a.astype(np.int16) + n, 0, 255).astype(np.uint8)
        n = np.random.default_rng(11).integers(-60, 61, a.shape, dtype=np.int16)
        big = np.clip(a.astype(np.int16) + n, 0, 255).astype(np.uint8)
        assert metrics_ref.psnr_simple(a, small) > metrics_ref.psnr_simple(a, big)
        assert metrics_ref.ssim_simple(a, small) > metrics_ref.ssim_simple(a, big)


class TestCv2Identities:  # the integer forms the kernels use
    def test_gray(self):
        a = _rand((90, 160, 3), 12)
        assert np.array_equal(metrics_ref.rgb2gray_u8(a), cv2.cvtColor(a, cv2.COLOR_RGB2GRAY))

    @pytest.mark.parametrize("H,W", [(360, 640), (1080, 1920)])
    def test_quarter_resize_is_centre_average(self, H, W):
        g = _rand((H, W), 13).astype(np.int32)
        want = cv2.resize(g.astype(np.uint8), None, fx=0.25, fy=0.25)
        got = (g[1::4, 1::4] + g[1::4, 2::4] + g[2::4, 1::4] + g[2::4, 2::4] + 2) >> 2
        assert np.array_equal(got.astype(np.uint8), want)

    @pytest.mark.parametrize("H,W", [(360, 640), (1080, 1920), (1440, 2560), (2160, 3840)])
    def test_hud_thumbnail_is_centre_average(self, H, W):
        g = _rand((H, W), 14).astype(np.int32)
        want = cv2.resize(g.astype(np.uint8), (320, 180))
        r = H // 180
        o = r // 2 - 1
        got = (g[o::r, o::r] + g[o::r, o + 1::r] + g[o + 1::r, o::r] + g[o + 1::r, o + 1::r] + 2) >> 2
        assert np.array_equal(got.astype(np.uint8), want)

    def test_laplacian_is_reflect101_stencil(self):
        g = _rand((40, 50), 15)
        p = np.pad(g.astype(np.int64), 1, mode="reflect")
        want = p[:-2, 1:-1] + p[2:, 1:-1] + p[1:-1, :-2] + p[1:-1, 2:] - 4 * p[1:-1, 1:-1]
        assert np.array_equal(cv2.Laplacian(g, cv2.CV_64F), want.astype(np.float64))
