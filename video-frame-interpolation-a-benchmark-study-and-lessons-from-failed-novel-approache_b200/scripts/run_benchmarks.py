"""Model registry + triplet benchmark loop, mirroring the reference's scripts/run_benchmarks.py for the
hot-path models: get_available_models() (:67-105) keeps the keys 'rife' and 'adaptive', zero-arg
constructible (:369); run_model_benchmark (:108-250) keeps its result keys for the in-memory case;
run_clip_benchmark is the same loop over a preprocessed clip DIRECTORY (load_clip_data's dict, :44-64, :108-250) with the
PNG decode prefetched on host threads (data_io.PrefetchReader) and both metrics of a triplet from one kernel pass."""
from dataclasses import asdict
from pathlib import Path
from typing import Dict, List, Optional

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


def _aggregate(values) -> dict:
    values = [v for v in values if v is not None]
    if not values:
        return {'mean': None}
    return {'mean': float(np.mean(values)), 'std': float(np.std(values)), 'min': float(np.min(values)),
            'max': float(np.max(values)), 'p50': float(np.percentile(values, 50)), 'p95': float(np.percentile(values, 95))}


def run_clip_benchmark(model, clip_data: dict, evaluator=None, max_triplets: Optional[int] = None,
                       save_outputs: bool = False, output_dir: Optional[Path] = None, threads: int = 8) -> dict:
    """The reference's run_model_benchmark(model, clip_data, evaluator, max_triplets, save_outputs, output_dir)
    (run_benchmarks.py:108-250): same arguments, same result dictionary (model, model_info, num_triplets, quality_summary,
    speed_summary, quality_per_triplet, speed_per_triplet), same per-triplet error tolerance.  Input and ground-truth PNGs
    are decoded ahead of the loop on `threads` host threads; quality = QualityEvaluator.evaluate (kernels)."""
    import cv2
    from ..data_io import PrefetchReader
    from ..evaluation.metrics import QualityEvaluator
    triplets = clip_data['triplets']
    if max_triplets:
        triplets = triplets[:max_triplets]
    input_dir, gt_dir, metadata = Path(clip_data['input_dir']), Path(clip_data['gt_dir']), clip_data['metadata']
    evaluator = evaluator or QualityEvaluator()
    num_intermediate = triplets[0]['num_intermediate'] if triplets else 3
    target_scale = metadata['spatial_scale']
    order, paths = [], []
    for tr in triplets:                                   # one flat decode stream: frame0, frame1, gt...
        order.append(2 + len(tr['gt_intermediate']))
        paths += [input_dir / tr['input_frame_0'], input_dir / tr['input_frame_1']] + [gt_dir / g for g in tr['gt_intermediate']]
    stream = iter(PrefetchReader(paths, threads=threads, ahead=4 * max(order, default=5)))
    quality_results, speed_results = [], []
    for tr, cnt in zip(triplets, order):
        try:
            frames = [next(stream)[1].numpy().copy() for _ in range(cnt)]
        except FileNotFoundError as e:
            print(f"    Error processing triplet {tr['triplet_id']}: {e}")
            break                                         # the flat stream is out of step after a missing file
        try:
            result = model.process_pair(frames[0], frames[1], num_intermediate=num_intermediate, target_scale=target_scale)
            pred, gt = result.frames[1:-1], frames[2:]
            if len(pred) != len(gt):
                print(f"    Warning: Predicted {len(pred)} frames, GT has {len(gt)}")
                continue
            quality_results.append(evaluator.evaluate(pred, gt).to_dict())
            speed_results.append({'inference_time_ms': result.inference_time_ms, 'vram_peak_mb': result.vram_peak_mb})
            if save_outputs and output_dir:
                d = Path(output_dir) / f"triplet_{tr['triplet_id']:04d}"
                d.mkdir(parents=True, exist_ok=True)
                for i, fr in enumerate(result.frames):
                    cv2.imwrite(str(d / f"frame_{i:02d}.png"), cv2.cvtColor(fr, cv2.COLOR_RGB2BGR))
        except Exception as e:
            print(f"    Error processing triplet {tr['triplet_id']}: {e}")
            continue
    if not quality_results:
        return {'error': 'No successful evaluations'}
    times = [r['inference_time_ms'] for r in speed_results]
    vram = [r['vram_peak_mb'] for r in speed_results]
    return {'model': model.info.name, 'model_info': asdict(model.info), 'num_triplets': len(quality_results),
            'quality_summary': {k: _aggregate([q.get(k) for q in quality_results]) for k in ('psnr', 'ssim', 'lpips')},
            'speed_summary': {'time_ms': {'mean': float(np.mean(times)), 'std': float(np.std(times)),
                                          'p50': float(np.percentile(times, 50)), 'p95': float(np.percentile(times, 95)),
                                          'p99': float(np.percentile(times, 99))},
                              'vram_mb': {'mean': float(np.mean(vram)), 'max': float(np.max(vram))},
                              'throughput_fps': 1000 / np.mean(times) * 5 if times else 0},
            'quality_per_triplet': quality_results, 'speed_per_triplet': speed_results}
