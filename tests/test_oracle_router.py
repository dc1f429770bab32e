"""oracle/router_ref.py pinned against the reference's own AdaptiveRouter (tests/golden/router.npz was produced
by running /root/reference models/novel/adaptive_pipeline.py:74-286; only skimage SSIM was shimmed)."""
import numpy as np
import pytest

import vfi_b200  # noqa: F401
from oracle import router_ref
from vfi_b200.synth import gen_clip

CODE = {"rife": 0, "vfimamba": 1, "scene_change": 2}


@pytest.fixture(scope="module")
def clip_and_gold(golden):
    g = golden("router.npz")
    n, H, W, seed = g["clip"].tolist()
    clip = gen_clip(n, H, W, seed=seed, cut_every=int(g["cut_every"]), fast=tuple(g["fast"].tolist()))
    return clip, g


def _rows(router, clip):
    rows, masks = [], []
    for i in range(len(clip) - 1):
        a = router.analyze(clip[i], clip[i + 1])
        rows.append([a.motion_mean, a.motion_max, a.motion_std, float(a.has_particles), float(a.is_scene_change),
                     a.hud_coverage, CODE[a.recommended_model], a.confidence])
        live = router.hud_mask is not None and not a.is_scene_change and len(router.frame_history) >= 5
        masks.append(np.packbits(router.hud_mask if live else np.zeros(clip[i].shape[:2], bool)))
    return np.array(rows), np.stack(masks)


def test_default_thresholds_bit_exact(clip_and_gold):
    clip, g = clip_and_gold
    rows, masks = _rows(router_ref.RouterRef(), clip)
    assert np.array_equal(rows, g["rows"])          # every float64 field identical
    assert np.array_equal(masks, g["masks"])        # every HUD mask bit identical
    assert set(rows[:, 6].astype(int)) == {0, 2}


def test_tight_thresholds_cover_all_branches(clip_and_gold):
    clip, g = clip_and_gold
    lo, hi, pt = g["tight"].tolist()
    rows, _ = _rows(router_ref.RouterRef(lo=lo, hi=hi, particle_thr=pt), clip)
    assert np.array_equal(rows, g["rows_tight"])
    assert set(rows[:, 6].astype(int)) == {0, 1, 2} and 0.7 in rows[:, 7] and 1.0 in rows[:, 3]


def test_pieces(clip_and_gold):
    clip, g = clip_and_gold
    assert [router_ref.laplacian_var(c) for c in clip] == g["laplacian_var"].tolist()
    import cv2
    for k, i in enumerate(g["motion_pairs"].tolist()):
        g0, g1 = (cv2.cvtColor(clip[j], cv2.COLOR_RGB2GRAY) for j in (i, i + 1))
        assert list(router_ref.flow_stats(router_ref.farneback(g0, g1))[:3]) == g["motion"][k].tolist()


def test_hud_composite():
    f0 = np.zeros((4, 4, 3), np.uint8)
    f1 = np.full((4, 4, 3), 9, np.uint8)
    m = np.zeros((4, 4), bool)
    m[0, :2] = True
    outs = router_ref.hud_composite([np.full((4, 4, 3), 5, np.uint8) for _ in range(3)], [0.25, 0.5, 0.75], f0, f1, m)
    assert outs[0][0, 0, 0] == 0 and outs[1][0, 0, 0] == 9 and outs[2][0, 1, 2] == 9 and outs[0][1, 1, 1] == 5


@pytest.mark.parametrize("H,W", [(480, 640), (720, 1280), (1000, 1777), (1080, 1920), (360, 640), (480, 854), (333, 517),
                                 (90, 160), (100, 100), (486, 482), (29, 31)])
def test_resize_restatement_matches_cv2(H, W):
    """oracle/resize_ref.py (what csrc/reduce.cu implements) == cv2.resize INTER_LINEAR on uint8, bit for bit, for the
    router's two call sites (adaptive_pipeline.py:149-155 fx=fy=0.25, :203-204 dsize=(320,180))."""
    import cv2
    from oracle.resize_ref import resize_linear_u8
    g = np.random.default_rng(H * 7 + W).integers(0, 256, (H, W), dtype=np.uint8)
    assert np.array_equal(resize_linear_u8(g, (320, 180)), cv2.resize(g, (320, 180)))
    assert np.array_equal(resize_linear_u8(g, fx=0.25, fy=0.25), cv2.resize(g, None, fx=0.25, fy=0.25))
