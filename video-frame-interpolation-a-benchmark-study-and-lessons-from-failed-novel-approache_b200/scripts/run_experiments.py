"""The RIFE / adaptive / metric pieces of the reference's experiment runner (scripts/run_experiments.py) on the B200
kernels -- the call sites that produced every published number of the study:

  RIFEModel.get          :252-271   singleton loader (train_log weights, or seeded random-init when opted in)
  calc_psnr / calc_ssim  :292-312   GLOBAL single-window PSNR / SSIM from five exact integer moments (vfi_moments_u8);
                                    the float64 formulas that follow are the reference's own expressions
  calc_motion            :359-362   Farneback(0.5, 1, 15, 1, 5, 1.1) mean |flow| (vfi_farneback)
  rife_interpolate       :365-386   BGR frames, pad to 128 with cv2.BORDER_REFLECT, Model.inference(t, scale), crop
                                    (vfi_ifnet_infer_ex with VFI_FLAG_BGR | VFI_FLAG_PAD_REFLECT_EDGE)
  adaptive_interpolate   :848-874   the evaluated "adaptive" method: 480x270 gray thumbnails -> calc_motion ->
                                    motion > thresh ? RIFE x3 : cv2.addWeighted x3
  estimate_motion_complexity / mcar_interpolate(_cached) :449-537   MCAR score (mean/20, std/15, Canny edge density) and its
                                    three tiers: addWeighted, cv2.remap flow warp + addWeighted (vfi_remap_flow_u8), RIFE

Frames are BGR uint8 (H,W,3), as the runner holds them (cv2.VideoCapture order).  No CPU fallback.
"""
import os
from typing import List, Optional, Tuple

import numpy as np
import torch

from .. import _abi, ifnet_params
from ..engine import Engine

FLAGS = _abi.FLAG_BGR | _abi.FLAG_PAD_REFLECT_EDGE


class RIFEModel:
    """Singleton RIFE loader (:252-271).  `get()` returns an object with the reference's inference surface reduced to what
    rife_interpolate needs: the engine with the IFNet weights loaded."""
    _instance = None
    precision = _abi.PRECISION_BF16

    @classmethod
    def get(cls, scale=0.5):
        if cls._instance is None:
            from ..models.sota.rife_wrapper import RIFEModel as Wrapper
            w = Wrapper()                      # same directory search / FileNotFoundError as the wrapper
            eng = Engine.get(None)
            eng.load_weights(w.state_dict())
            cls._instance = eng
            print("[RIFE] Model loaded on GPU")
        return cls._instance


def _moments(img1: np.ndarray, img2: np.ndarray):
    m = Engine.get(None).moments(np.ascontiguousarray(img1)[None], np.ascontiguousarray(img2)[None])[0].tolist()
    return m, img1.size


def calc_psnr(img1, img2) -> float:
    """:292-302 -- mse = mean((a-b)^2) is an exact integer sum / N; then the reference's float64 expression."""
    (sa, sb, saa, sbb, sab), n = _moments(img1, img2)
    mse = np.float64(saa + sbb - 2 * sab) / n
    if mse == 0:
        return float('inf')
    return float(20 * np.log10(255.0 / np.sqrt(mse)))


def ssim_from_moments(sa, sb, saa, sbb, sab, n) -> float:
    """:305-312 with the means / (co)variances formed from the exact moments in float64 (the reference sums the float64
    deviations; the two agree to ~1e-13 relative)."""
    C1, C2 = (0.01 * 255) ** 2, (0.03 * 255) ** 2
    mu1, mu2 = np.float64(sa) / n, np.float64(sb) / n
    # integer numerators: n*sum(x^2) - (sum x)^2 etc. are exact in Python ints
    sigma1_sq = np.float64(n * saa - sa * sa) / (np.float64(n) * n)
    sigma2_sq = np.float64(n * sbb - sb * sb) / (np.float64(n) * n)
    sigma12 = np.float64(n * sab - sa * sb) / (np.float64(n) * n)
    ssim = ((2 * mu1 * mu2 + C1) * (2 * sigma12 + C2)) / ((mu1 ** 2 + mu2 ** 2 + C1) * (sigma1_sq + sigma2_sq + C2))
    return float(max(0.0, min(1.0, ssim)))


def calc_ssim(img1, img2) -> float:
    (sa, sb, saa, sbb, sab), n = _moments(img1, img2)
    return ssim_from_moments(sa, sb, saa, sbb, sab, n)


def calc_motion(prev_gray: np.ndarray, curr_gray: np.ndarray) -> float:
    """:359-362 -- mean |Farneback(0.5, 1, 15, 1, 5, 1.1)| of two gray images."""
    st = Engine.get(None).farneback(prev_gray[None], curr_gray[None], pyr_scale=0.5, levels=1, winsize=15, iterations=1,
                                    poly_n=5, poly_sigma=1.1, want_flow=False, want_stats=True)
    return float(st[0, 0])


def rife_interpolate(prev: np.ndarray, curr: np.ndarray, t: float, scale: float = 0.5) -> np.ndarray:
    """:365-386 -- one interpolated BGR frame at timestep t."""
    return rife_interpolate_multi(prev, curr, [t], scale)[0]


def rife_interpolate_multi(prev, curr, ts, scale: float = 0.5) -> List[np.ndarray]:
    """The runner calls rife_interpolate three times per pair (i/4, :861,:883); one call here shares the Head features."""
    eng = RIFEModel.get()
    out = eng.infer(np.ascontiguousarray(prev)[None], np.ascontiguousarray(curr)[None], list(ts), scale,
                    RIFEModel.precision, flags=FLAGS)
    host = out[0].cpu().numpy()
    return [host[i] for i in range(len(ts))]


def adaptive_interpolate(prev, curr, motion_thresh: float, rife_scale: float = 0.5,
                         prev_gray=None, curr_gray=None) -> Tuple[List[np.ndarray], bool, float]:
    """:848-874 -- the three intermediate frames of one pair, whether RIFE ran, and the motion value routed on.
    prev_gray / curr_gray (cv2 BGR2GRAY of the frames) are computed on the device when not given."""
    eng = Engine.get(None)
    if prev_gray is None:
        th = eng.resize_gray(np.stack([prev, curr]), (480, 270), bgr=True)
    else:
        th = eng.resize_gray(np.stack([prev_gray, curr_gray]), (480, 270))
    st = eng.farneback(th[:1], th[1:], pyr_scale=0.5, levels=1, winsize=15, iterations=1, poly_n=5, poly_sigma=1.1,
                       want_flow=False, want_stats=True)
    motion = float(st[0, 0])
    ts = [i / 4 for i in range(1, 4)]
    if motion > motion_thresh:
        return rife_interpolate_multi(prev, curr, ts, rife_scale), True, motion
    a, b = eng._u8(prev), eng._u8(curr)
    return [eng.add_weighted(a, 1 - t, b, t).cpu().numpy() for t in ts], False, motion


def estimate_motion_complexity(prev_gray, curr_gray):
    """:449-476 -- (complexity in [0,1], flow).  Farneback + |flow| mean/std on the device; Canny edge density on the
    device (vfi_canny_density)."""
    eng = Engine.get(None)
    flow, st = eng.farneback(eng._u8(prev_gray)[None], eng._u8(curr_gray)[None], pyr_scale=0.5, levels=1, winsize=15, iterations=1,
                             poly_n=5, poly_sigma=1.1, want_flow=True, want_stats=True)
    mean, _, std = st[0].tolist()
    mean_motion = min(1.0, mean / 20.0)
    motion_var = min(1.0, std / 15.0)
    edge_density = float(eng.canny_density(eng._u8(curr_gray)[None], 100, 200)[0]) * 2.0
    complexity = min(1.0, 0.5 * mean_motion + 0.3 * motion_var + 0.2 * edge_density)
    return complexity, flow[0]


def mcar_tier(complexity: float, threshold_low: float = 0.3, threshold_high: float = 0.7) -> str:
    """:479-520 -- which interpolator MCAR picks for a complexity score."""
    if complexity < threshold_low:
        return 'linear'
    if complexity < threshold_high:
        return 'flow'
    return 'rife'


def mcar_interpolate_cached(prev, curr, t, complexity, flow, rife_scale=0.5, threshold_low=0.3, threshold_high=0.7):
    """:513-537 -- one MCAR frame from a precomputed complexity and flow: (frame, tier).  flow: the device tensor
    estimate_motion_complexity returned (or a host array)."""
    eng = Engine.get(None)
    tier = mcar_tier(complexity, threshold_low, threshold_high)
    if tier == 'linear':
        return eng.add_weighted(prev, 1 - t, curr, t).cpu().numpy(), tier
    if tier == 'flow':
        warped = eng.remap_flow(np.ascontiguousarray(prev), flow, t)
        return eng.add_weighted(warped, 1 - t, curr, t).cpu().numpy(), tier
    return rife_interpolate(prev, curr, t, rife_scale), tier


def mcar_interpolate(prev, curr, t, rife_scale=0.5, threshold_low=0.3, threshold_high=0.7):
    """:479-510 -- (frame, tier, complexity); the grays are cv2's BGR2GRAY computed on the device."""
    eng = Engine.get(None)
    g = eng.resize_gray(np.stack([prev, curr]), (prev.shape[1], prev.shape[0]), bgr=True)
    complexity, flow = estimate_motion_complexity(g[0], g[1])
    frame, tier = mcar_interpolate_cached(prev, curr, t, complexity, flow, rife_scale, threshold_low, threshold_high)
    return frame, tier, complexity
