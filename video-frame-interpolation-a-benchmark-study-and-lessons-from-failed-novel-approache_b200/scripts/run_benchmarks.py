"""Model registry + triplet benchmark loop, mirroring the reference's scripts/run_benchmarks.py for the
hot-path models: get_available_models() (:67-105) keeps the keys 'rife' and 'adaptive', zero-arg
constructible (:369); run_model_benchmark (:108-250) keeps its result keys for the in-memory case."""
from typing import Dict, List

import numpy as np


def get_available_models() -> Dict[str, type]:
    models = {}
    try:
        from ..models.sota.rife_wrapper import RIFEModel
        models['rife'] = RIFEModel
    except ImportError as e:
        print(f"Warning: Could not import RIFE: {e}")
    try:
        from ..models.novel.adaptive_pipeline import AdaptivePipeline
        models['adaptive'] = AdaptivePipeline
    except ImportError as e:
        print(f"Warning: Could not import AdaptivePipeline: {e}")
    return models


def run_model_benchmark(model, triplets: List[dict], num_intermediate: int = 3, target_scale: float = 1.0,
                        evaluate: bool = True) -> dict:
    """triplets: [{'frame0': u8 HWC, 'frame1': u8 HWC, 'gt': [u8 HWC]*num_intermediate (optional)}].
    Returns the reference's aggregate keys (timing percentiles, throughput_fps, psnr/ssim means)."""
    from ..evaluation.metrics import compute_psnr_simple, compute_ssim_simple
    times, psnrs, ssims, vram = [], [], [], []
    for tr in triplets:
        try:
            res = model.process_pair(tr['frame0'], tr['frame1'], num_intermediate, target_scale)
        except Exception as e:  # same per-triplet tolerance as the reference (:143,:196-198)
            print(f"Error processing triplet: {e}")
            continue
        times.append(res.inference_time_ms)
        vram.append(res.vram_peak_mb)
        if evaluate and tr.get('gt') is not None:
            for pred, gt in zip(res.frames[1:-1], tr['gt']):
                psnrs.append(compute_psnr_simple(pred, gt))
                ssims.append(compute_ssim_simple(pred, gt))
    if not times:
        return {'error': 'no triplets processed'}
    t = np.asarray(times)
    out = {'model': model.info.name, 'num_triplets': len(times), 'time_mean_ms': float(t.mean()),
           'time_std_ms': float(t.std()), 'time_p50_ms': float(np.percentile(t, 50)),
           'time_p95_ms': float(np.percentile(t, 95)), 'time_p99_ms': float(np.percentile(t, 99)),
           'throughput_fps': float(1000.0 / t.mean() * (num_intermediate + 2)), 'vram_peak_mb': float(max(vram))}
    if psnrs:
        finite = [p for p in psnrs if np.isfinite(p)]
        out.update(psnr_mean=float(np.mean(finite)) if finite else float('inf'), ssim_mean=float(np.mean(ssims)))
    return out
