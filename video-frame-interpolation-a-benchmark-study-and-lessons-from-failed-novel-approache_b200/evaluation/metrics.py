"""PSNR / SSIM of the reference's evaluation/metrics.py, computed by the B200 reduction kernels.

compute_psnr_simple (reference :299-315) and compute_ssim_simple (:318-335) keep their signatures and
results: PSNR from the exact integer sum of squared differences (so the float64 arithmetic that follows
is numpy's own and the value is bit-identical), SSIM = skimage structural_similarity on the cv2 gray
images (7x7 uniform window) to ~1e-12.  QualityEvaluator / QualityResults (:88-258) keep their surface;
psnr/ssim are served by the kernels, the pyiqa-only metrics (lpips, ms_ssim, flolpips) are out of scope
and reported as None / NaN.  Batched variants (psnr_batch / ssim_batch) are the throughput entry points.
"""
import dataclasses
from typing import Dict, List, Optional

import cv2
import numpy as np
import torch

from ..engine import Engine


def _match(pred: np.ndarray, gt: np.ndarray) -> np.ndarray:
    if pred.shape != gt.shape:
        pred = cv2.resize(pred, (gt.shape[1], gt.shape[0]))  # same default INTER_LINEAR as the reference
    return pred


def _psnr_from_sse(sse: int, n: int) -> float:
    mse = np.float64(sse) / n          # == np.mean of the exact integer squares
    if mse == 0:
        return float('inf')
    return float(10 * np.log10(255.0 ** 2 / mse))


def compute_psnr_simple(pred: np.ndarray, gt: np.ndarray) -> float:
    pred = _match(pred, gt)
    sse, _ = Engine.get().psnr_sse(pred[None], gt[None])
    return _psnr_from_sse(int(sse[0]), pred.size)


def compute_ssim_simple(pred: np.ndarray, gt: np.ndarray) -> float:
    pred = _match(pred, gt)
    if pred.ndim == 2:                 # already gray: replicate so the integer gray conversion is the identity
        pred, gt = (np.repeat(a[..., None], 3, axis=2) for a in (pred, gt))
    return float(Engine.get().ssim(pred[None], gt[None], downscale=1)[0])


def psnr_batch(pred, gt, device=None) -> np.ndarray:
    """(n,H,W,3) uint8 numpy / CUDA tensors -> n PSNR values (float64, inf for identical frames)."""
    sse, _ = Engine.get(device).psnr_sse(pred, gt)
    n = int(np.prod(pred.shape[1:]))
    return np.array([_psnr_from_sse(int(s), n) for s in sse.tolist()], dtype=np.float64)


def ssim_batch(pred, gt, device=None) -> np.ndarray:
    return Engine.get(device).ssim(pred, gt, downscale=1).cpu().numpy()


def psnr_ssim_batch(pred, gt, device=None):
    """Both metrics of (n,H,W,3) uint8 frame pairs from one pass over the frames (vfi_psnr_ssim_u8): the values of
    psnr_batch and ssim_batch.  Returns (psnr float64 (n,), ssim float64 (n,), sse int64 (n,)) as numpy arrays."""
    sse, _, ss = Engine.get(device).psnr_ssim(pred, gt)
    n = int(np.prod(pred.shape[1:]))
    sse = sse.cpu().numpy()
    return np.array([_psnr_from_sse(int(s), n) for s in sse.tolist()], dtype=np.float64), ss.cpu().numpy(), sse


def evaluate_sharded(pred, gt, rank: int = 0, world: int = 1, group=None, chunk: int = 64, device=None) -> dict:
    """Extra (not in the reference): PSNR / SSIM means of n frame pairs sharded by frame index over `world` ranks
    (BASELINE.json config 5; the means the reference forms at evaluation/metrics.py:208-258 from its per-frame values).
    pred, gt: (n,H,W,3) uint8 -- numpy or CUDA tensors holding the WHOLE set (each rank touches only its contiguous
    range; evaluate_range takes producers instead when no rank can hold the whole set).  Each rank runs the fused
    PSNR+SSIM kernel (vfi_psnr_ssim_u8) over its range `chunk` pairs at a time, and ONE all-reduce(SUM) of
    {sum psnr over finite values, sum ssim, n, n identical (psnr = inf), sum of integer SSE} ends the job (float64
    vector; the integer entries stay exact below 2^53).  Returns the whole-set means and sums on every rank."""
    from .. import sharding
    e = Engine.get(device)
    lo, hi = sharding.shard_range(pred.shape[0], rank, world)
    part = evaluate_range(lambda a, b: pred[a:b], lambda a, b: gt[a:b], lo, hi, chunk, e)
    return reduce_sums(part, (lo, hi), group, e.device)


def reduce_sums(part, rng=None, group=None, device=None) -> dict:
    """The job's one collective: all-reduce(SUM) of a rank's evaluate_range() vector, then the means."""
    from .. import sharding
    vec = torch.tensor(part, dtype=torch.float64, device=device if device is not None else 'cpu')
    return _finish_sums(sharding.reduce_counters(vec, group).tolist(), rng)


def evaluate_range(pred_fn, gt_fn, lo: int, hi: int, chunk: int = 64, engine=None, device=None):
    """Sums over frame pairs [lo, hi): [sum psnr (finite), sum ssim, n, n_inf, sum sse] as Python floats/ints.
    The per-frame PSNR is formed on the host from the exact integer SSE with the reference's own float64 expression
    (evaluation/metrics.py:311-315), so the sum is over bit-identical per-frame values."""
    e = engine or Engine.get(device)
    s_psnr = s_ssim = 0.0
    n_tot = n_inf = sse_tot = 0
    pending = []
    for a in range(lo, hi, chunk):
        b = min(hi, a + chunk)
        p, g = pred_fn(a, b), gt_fn(a, b)
        sse, _, ss = e.psnr_ssim(p, g)
        pending.append((sse, ss, int(np.prod(p.shape[1:]))))
    for sse, ss, elems in pending:                    # one device->host read per chunk, after everything is queued
        sse = sse.tolist()
        for v in sse:
            ps = _psnr_from_sse(int(v), elems)
            if np.isinf(ps):
                n_inf += 1
            else:
                s_psnr += ps
            sse_tot += int(v)
        s_ssim += float(ss.sum(dtype=torch.float64))
        n_tot += len(sse)
    assert sse_tot < 2 ** 53
    return [s_psnr, s_ssim, float(n_tot), float(n_inf), float(sse_tot)]


def _finish_sums(vec, rng=None) -> dict:
    s_psnr, s_ssim, n, n_inf, sse = vec
    n, n_inf = int(n), int(n_inf)
    return {'range': rng, 'n': n, 'n_identical': n_inf, 'sse_sum': int(sse), 'psnr_sum': s_psnr, 'ssim_sum': s_ssim,
            'psnr': float('inf') if n_inf else (s_psnr / n if n else float('nan')),     # np.mean with an inf entry is inf
            'ssim': s_ssim / n if n else float('nan')}


@dataclasses.dataclass
class QualityResults:
    psnr: float
    ssim: float
    lpips: float
    ms_ssim: Optional[float] = None
    flolpips: Optional[float] = None
    psnr_per_frame: Optional[List[float]] = None
    lpips_per_frame: Optional[List[float]] = None

    def to_dict(self) -> dict:
        return {'psnr': self.psnr, 'ssim': self.ssim, 'lpips': self.lpips, 'ms_ssim': self.ms_ssim,
                'flolpips': self.flolpips}


class QualityEvaluator:
    """Reference surface (evaluation/metrics.py:111-258).  psnr / ssim come from the kernels (the
    definitions of compute_*_simple, which is also what the reference driver falls back to at
    scripts/run_benchmarks.py:174-178 when pyiqa is unavailable); lpips / ms_ssim / flolpips need pyiqa's
    networks and are not provided."""

    def __init__(self, device: str = 'cuda'):
        self.device = device
        self._initialized = False

    def _ensure_initialized(self):
        if not self._initialized:
            Engine.get()
            self._initialized = True

    def evaluate_pair(self, pred: np.ndarray, gt: np.ndarray) -> Dict[str, float]:
        self._ensure_initialized()
        return {'psnr': compute_psnr_simple(pred, gt), 'ssim': compute_ssim_simple(pred, gt), 'lpips': float('nan')}

    def evaluate(self, pred_frames: List[np.ndarray], gt_frames: List[np.ndarray]) -> QualityResults:
        self._ensure_initialized()
        if len(pred_frames) != len(gt_frames):
            raise ValueError(f"Number of predicted frames ({len(pred_frames)}) "
                             f"doesn't match ground truth ({len(gt_frames)})")
        same = len({f.shape for f in pred_frames} | {f.shape for f in gt_frames}) == 1
        if same and pred_frames:
            p, g = np.stack(pred_frames), np.stack(gt_frames)
            ps, ss, _ = psnr_ssim_batch(p, g)
            ps, ss = ps.tolist(), ss.tolist()
        else:
            ps = [compute_psnr_simple(a, b) for a, b in zip(pred_frames, gt_frames)]
            ss = [compute_ssim_simple(a, b) for a, b in zip(pred_frames, gt_frames)]
        return QualityResults(psnr=float(np.mean(ps)), ssim=float(np.mean(ss)), lpips=float('nan'),
                              psnr_per_frame=ps, lpips_per_frame=[float('nan')] * len(ps))
