"""oracle/farneback_ref.py (the numpy restatement of OpenCV's Farneback that csrc/farneback.cu implements) pinned against
cv2.calcOpticalFlowFarneback itself, with the reference's two parameter sets (models/novel/adaptive_pipeline.py:119-128;
scripts/run_experiments.py:359-362)."""
import cv2
import numpy as np
import pytest

import vfi_b200  # noqa: F401
from oracle import farneback_ref
from vfi_b200.synth import gen_clip


@pytest.mark.parametrize("H,W,seed", [(135, 240, 1), (203, 317, 3), (270, 480, 2)])
def test_restatement_matches_cv2(H, W, seed):
    clip = gen_clip(2, H, W, seed=seed, cut_every=0, fast=(0, 2))
    for i in range(2):
        g0, g1 = (cv2.cvtColor(clip[i + k], cv2.COLOR_RGB2GRAY) for k in (0, 1))
        ref = cv2.calcOpticalFlowFarneback(g0, g1, None, 0.5, 3, 15, 3, 5, 1.2, 0)
        got = farneback_ref.farneback(g0, g1)
        assert got.shape == ref.shape and got.dtype == np.float32
        assert float(np.abs(got - ref).max()) <= 1e-3 and float(np.abs(got - ref).mean()) <= 1e-5
        assert float(np.sqrt((ref ** 2).sum(-1)).max()) > 3.0          # the clip really moves


def test_restatement_matches_cv2_single_level_router():
    """The evaluated motion-threshold router's call (run_experiments.py:359-362): levels 1, iterations 1, sigma 1.1 on 480x270."""
    clip = gen_clip(1, 270, 480, seed=5, cut_every=0, fast=None)
    g0, g1 = (cv2.cvtColor(clip[k], cv2.COLOR_RGB2GRAY) for k in (0, 1))
    ref = cv2.calcOpticalFlowFarneback(g0, g1, None, 0.5, 1, 15, 1, 5, 1.1, 0)
    got = farneback_ref.farneback(g0, g1, levels=1, iterations=1, poly_sigma=1.1)
    assert float(np.abs(got - ref).max()) <= 1e-3


def test_pyramid_geometry():
    assert farneback_ref.pyramid_levels(1920, 1080) == 3 and farneback_ref.pyramid_levels(100, 100) == 1
    assert farneback_ref.pyramid_levels(60, 60) == 0
    assert [farneback_ref.level_geometry(1920, 1080, k)[2:] for k in range(4)] == [
        (3, 1920, 1080), (3, 960, 540), (9, 480, 270), (19, 240, 135)]
