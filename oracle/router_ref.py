"""CPU oracle: AdaptiveRouter / AdaptivePipeline routing semantics (cv2 + numpy + oracle SSIM).

TEST INFRASTRUCTURE ONLY (see oracle/ifnet_ref.py header for who may import it).

Follows models/novel/adaptive_pipeline.py:74-286 (router) and :369-411 (interpolate's
routing + HUD composite).  cv2 (opencv-python-headless 4.13.0 here; requirements.txt:15
asks >= 4.8) is called directly for gray / resize / Laplacian / morphology / Farneback,
exactly as the reference does; skimage SSIM is the restatement in oracle/metrics_ref.py.
PARITY UNPINNED: the reference has no test for any AdaptiveRouter output; where the
reference itself can run here (everything except skimage SSIM) tests/golden/make_golden.py
runs the reference class and this oracle side by side.
"""
from collections import deque
from dataclasses import dataclass

import cv2
import numpy as np

from .metrics_ref import ssim_gray_u8

ROUTE_RIFE, ROUTE_VFIMAMBA, ROUTE_SCENE = 0, 1, 2
ROUTE_NAMES = ("rife", "vfimamba", "scene_change")


@dataclass
class Analysis:
    motion_mean: float
    motion_max: float
    motion_std: float
    has_particles: bool
    is_scene_change: bool
    hud_coverage: float
    recommended_model: str
    confidence: float
    ssim_score: float = 0.0
    particle_score: float = 0.0
    laplacian_var: float = 0.0


def farneback(gray0, gray1):
    """adaptive_pipeline.py:119-128 parameters."""
    return cv2.calcOpticalFlowFarneback(gray0, gray1, None, pyr_scale=0.5, levels=3, winsize=15,
                                        iterations=3, poly_n=5, poly_sigma=1.2, flags=0)


def flow_stats(flow):
    """adaptive_pipeline.py:130-138: f32 magnitude, numpy mean / max / std."""
    mag = np.sqrt(flow[..., 0] ** 2 + flow[..., 1] ** 2)
    return float(np.mean(mag)), float(np.max(mag)), float(np.std(mag)), mag


def scene_ssim(frame0, frame1):
    """adaptive_pipeline.py:140-160 (gray -> x0.25 INTER_LINEAR -> SSIM)."""
    g0 = cv2.cvtColor(frame0, cv2.COLOR_RGB2GRAY)
    g1 = cv2.cvtColor(frame1, cv2.COLOR_RGB2GRAY)
    s0 = cv2.resize(g0, None, fx=0.25, fy=0.25)
    s1 = cv2.resize(g1, None, fx=0.25, fy=0.25)
    return ssim_gray_u8(s0, s1)


def laplacian_var(frame):
    """adaptive_pipeline.py:180-183."""
    gray = cv2.cvtColor(frame, cv2.COLOR_RGB2GRAY)
    return float(cv2.Laplacian(gray, cv2.CV_64F).var())


def decide(has_particles, particle_score, mmean, mmax, lo, hi):
    """adaptive_pipeline.py:266-275."""
    if has_particles or mmax > hi:
        return "vfimamba", min(particle_score + mmax / 50.0, 1.0)
    if mmean < lo:
        return "rife", 1.0 - mmean / lo
    return "rife", 0.7


class RouterRef:
    def __init__(self, lo=5.0, hi=25.0, scene_thr=0.65, particle_thr=0.4, hud_var_thr=10.0, hud_hist=10,
                 flow_fn=None):
        self.lo, self.hi = lo, hi
        self.scene_thr, self.particle_thr = scene_thr, particle_thr
        self.hud_var_thr = hud_var_thr
        self.frame_history = deque(maxlen=hud_hist)
        self.hud_mask = None
        self.flow_fn = flow_fn or farneback

    def detect_hud(self, frame):
        """adaptive_pipeline.py:193-235."""
        gray = cv2.cvtColor(frame, cv2.COLOR_RGB2GRAY)
        self.frame_history.append(cv2.resize(gray, (320, 180)))
        if len(self.frame_history) < 5:
            return np.zeros(frame.shape[:2], dtype=bool), 0.0
        var = np.var(np.array(list(self.frame_history)[-5:]), axis=0)
        small = (var < self.hud_var_thr).astype(np.uint8)
        m = cv2.resize(small, (frame.shape[1], frame.shape[0]), interpolation=cv2.INTER_NEAREST)
        k = np.ones((5, 5), np.uint8)
        m = cv2.morphologyEx(m, cv2.MORPH_CLOSE, k)
        m = cv2.morphologyEx(m, cv2.MORPH_OPEN, k)
        self.hud_mask = m.astype(bool)
        return self.hud_mask, float(self.hud_mask.mean())

    def analyze(self, frame0, frame1):
        """adaptive_pipeline.py:237-286."""
        score = scene_ssim(frame0, frame1)
        if score < self.scene_thr:
            return Analysis(0, 0, 0, False, True, 0, "scene_change", 1.0 - score, ssim_score=score)
        g0 = cv2.cvtColor(frame0, cv2.COLOR_RGB2GRAY)
        g1 = cv2.cvtColor(frame1, cv2.COLOR_RGB2GRAY)
        mmean, mmax, mstd, _ = flow_stats(self.flow_fn(g0, g1))
        lv = laplacian_var(frame0)
        pscore = float(np.sqrt(min(mstd / 20.0, 1.0) * min(lv / 500.0, 1.0)))
        has_p = pscore > self.particle_thr
        _, cov = self.detect_hud(frame0)
        model, conf = decide(has_p, pscore, mmean, mmax, self.lo, self.hi)
        return Analysis(mmean, mmax, mstd, bool(has_p), False, cov, model, float(conf),
                        ssim_score=score, particle_score=pscore, laplacian_var=lv)


def hud_composite(frames, timestamps, frame0, frame1, hud_mask):
    """adaptive_pipeline.py:399-409."""
    for fr, t in zip(frames, timestamps):
        src = frame0 if t < 0.5 else frame1
        fr[hud_mask] = src[hud_mask]
    return frames
