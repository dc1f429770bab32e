"""On-disk clip formats of the reference and a prefetching reader that keeps the GPU fed.

Triplet layout (scripts/preprocess_video.py:349-470, read by scripts/run_benchmarks.py:36-64):
    <clip>/metadata.json, <clip>/triplets.json, <clip>/input_1080p30/frames/frame_*.png, <clip>/ground_truth/frames/frame_*.png
Interval layout (scripts/run_experiments.py:52-127; data/clips/<clip>/intervals/<interval>/):
    meta.json, keyframes/kf_*.png, midpoints/gt_*.png, motion/flow_stats.json
PNG decode stays on the host (cv2, as the reference; there is no PNG path on the GPU's decode engines) but runs on a pool
of threads AHEAD of the consumer into pinned buffers, so that the hot loop sees uint8 frames already in page-locked
memory: the reference's loop decodes, runs and evaluates serially per triplet (run_benchmarks.py:142-178).
"""
import json
import os
from collections import deque
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import Callable, Iterable, Iterator, List, Optional, Sequence

import cv2
import numpy as np
import torch


# ---- the reference's readers ---------------------------------------------------------------------------------------
def load_image(path) -> np.ndarray:
    """RGB uint8 (run_benchmarks.py:36-41)."""
    img = cv2.imread(str(path))
    if img is None:
        raise FileNotFoundError(f"Could not load image: {path}")
    return cv2.cvtColor(img, cv2.COLOR_BGR2RGB)


def load_clip_data(clip_dir) -> dict:
    """run_benchmarks.py:44-64 -- same keys, same FileNotFoundError messages."""
    clip_dir = Path(clip_dir)
    metadata_path, triplets_path = clip_dir / 'metadata.json', clip_dir / 'triplets.json'
    if not metadata_path.exists():
        raise FileNotFoundError(f"Metadata not found: {metadata_path}")
    if not triplets_path.exists():
        raise FileNotFoundError(f"Triplets not found: {triplets_path}")
    with open(metadata_path) as f:
        metadata = json.load(f)
    with open(triplets_path) as f:
        triplets = json.load(f)
    return {'metadata': metadata, 'triplets': triplets, 'input_dir': clip_dir / 'input_1080p30' / 'frames',
            'gt_dir': clip_dir / 'ground_truth' / 'frames'}


def create_triplet_index(input_names: Sequence[str], gt_names: Sequence[str], input_fps: float, target_fps: float) -> List[dict]:
    """preprocess_video.py:349-424 -- the triplet index of a 30 -> 120 fps clip (same fields, same rounding)."""
    if not input_names:
        raise ValueError("No input frames found")
    if not gt_names:
        raise ValueError("No ground truth frames found")
    temporal_scale = target_fps / input_fps
    triplets = []
    for i in range(len(input_names) - 1):
        gt_start, gt_end = int(round(i * temporal_scale)), int(round((i + 1) * temporal_scale))
        mid = list(range(gt_start + 1, gt_end))
        if not mid or gt_end >= len(gt_names):
            continue
        triplets.append({
            'triplet_id': len(triplets), 'input_frame_0': input_names[i], 'input_frame_1': input_names[i + 1],
            'input_idx_0': i, 'input_idx_1': i + 1, 'gt_frame_start': gt_names[gt_start], 'gt_frame_end': gt_names[gt_end],
            'gt_start_idx': gt_start, 'gt_end_idx': gt_end,
            'gt_intermediate': [gt_names[k] for k in mid if k < len(gt_names)],
            'gt_intermediate_indices': [k for k in mid if k < len(gt_names)],
            'temporal_positions': [(k - gt_start) / (gt_end - gt_start) for k in mid], 'num_intermediate': len(mid)})
    return triplets


def load_interval(clips_dir, clip_id: str, interval_id: str) -> Optional[dict]:
    """run_experiments.py:52-106 -- BGR frames, as cv2.imread returns them (the experiment runner works in BGR)."""
    d = Path(clips_dir) / clip_id / 'intervals' / interval_id
    if not d.exists() or not (d / 'meta.json').exists():
        return None
    with open(d / 'meta.json') as f:
        meta = json.load(f)
    read = lambda sub, pat: [fr for fr in (cv2.imread(str(p)) for p in sorted((d / sub).glob(pat))) if fr is not None]
    stats = {}
    if (d / 'motion' / 'flow_stats.json').exists():
        with open(d / 'motion' / 'flow_stats.json') as f:
            stats = json.load(f)
    return {'keyframes': read('keyframes', 'kf_*.png'), 'midpoints': read('midpoints', 'gt_*.png'), 'meta': meta,
            'motion_stats': stats, 'clip_id': clip_id, 'interval_id': interval_id}


def list_available_intervals(clips_dir, clip_id: str) -> List[dict]:
    """run_experiments.py:109-123."""
    d = Path(clips_dir) / clip_id / 'intervals'
    if not d.exists():
        return []
    out = []
    for p in sorted(d.glob('interval_*')):
        if (p / 'meta.json').exists():
            with open(p / 'meta.json') as f:
                out.append(json.load(f))
    return out


# ---- writers (synthetic clips in the reference's layouts: tests, bench fixtures) ---------------------------------------
def write_clip(clip_dir, gt_frames: np.ndarray, temporal_scale: int = 4, spatial_scale: float = 1.0, input_fps: float = 30.0) -> dict:
    """Write RGB frames (n,H,W,3) as a triplet clip: every `temporal_scale`-th frame is an input frame, all are ground truth."""
    clip_dir = Path(clip_dir)
    in_dir, gt_dir = clip_dir / 'input_1080p30' / 'frames', clip_dir / 'ground_truth' / 'frames'
    in_dir.mkdir(parents=True, exist_ok=True)
    gt_dir.mkdir(parents=True, exist_ok=True)
    gt_names, in_names = [], []
    for i, fr in enumerate(gt_frames):
        name = f'frame_{i:06d}.png'
        cv2.imwrite(str(gt_dir / name), cv2.cvtColor(fr, cv2.COLOR_RGB2BGR))
        gt_names.append(name)
        if i % temporal_scale == 0:
            iname = f'frame_{i // temporal_scale:06d}.png'
            cv2.imwrite(str(in_dir / iname), cv2.cvtColor(fr, cv2.COLOR_RGB2BGR))
            in_names.append(iname)
    triplets = create_triplet_index(in_names, gt_names, input_fps, input_fps * temporal_scale)
    H, W = gt_frames.shape[1:3]
    metadata = {'source_path': 'synthetic', 'source_resolution': [W, H], 'input_resolution': [W, H], 'input_fps': input_fps,
                'target_resolution': [int(W * spatial_scale), int(H * spatial_scale)], 'target_fps': input_fps * temporal_scale,
                'temporal_scale': float(temporal_scale), 'spatial_scale': spatial_scale, 'num_input_frames': len(in_names),
                'num_gt_frames': len(gt_names), 'num_triplets': len(triplets)}
    with open(clip_dir / 'metadata.json', 'w') as f:
        json.dump(metadata, f, indent=2)
    with open(clip_dir / 'triplets.json', 'w') as f:
        json.dump(triplets, f, indent=2)
    return {'metadata': metadata, 'triplets': triplets}


def write_interval(clips_dir, clip_id: str, interval_id: str, frames_bgr: np.ndarray, motion_stats: Optional[dict] = None) -> Path:
    """Even frames -> keyframes/kf_*.png, odd frames -> midpoints/gt_*.png (extract_intervals.py layout)."""
    d = Path(clips_dir) / clip_id / 'intervals' / interval_id
    (d / 'keyframes').mkdir(parents=True, exist_ok=True)
    (d / 'midpoints').mkdir(parents=True, exist_ok=True)
    kf = gt = 0
    for i, fr in enumerate(frames_bgr):
        if i % 2 == 0:
            cv2.imwrite(str(d / 'keyframes' / f'kf_{kf:05d}.png'), fr)
            kf += 1
        else:
            cv2.imwrite(str(d / 'midpoints' / f'gt_{gt:05d}.png'), fr)
            gt += 1
    meta = {'interval_id': interval_id, 'clip_id': clip_id, 'start_frame': 0, 'end_frame': len(frames_bgr) - 1,
            'keyframe_count': kf, 'midpoint_count': gt, 'resolution': f'{frames_bgr.shape[2]}x{frames_bgr.shape[1]}',
            'motion_stats': motion_stats or {}}
    with open(d / 'meta.json', 'w') as f:
        json.dump(meta, f, indent=2)
    if motion_stats:
        (d / 'motion').mkdir(exist_ok=True)
        with open(d / 'motion' / 'flow_stats.json', 'w') as f:
            json.dump(motion_stats, f)
    return d


# ---- prefetching reader ----------------------------------------------------------------------------------------------
class PrefetchReader:
    """Decode a sequence of image files on `threads` host threads, `ahead` files ahead of the consumer, each into its own
    pinned uint8 buffer (so the consumer's H2D copy is asynchronous).  Iterating yields (index, pinned tensor (H,W,3) RGB)
    in order.  Buffers are recycled `ahead + 2` items after they were handed out."""

    def __init__(self, paths: Iterable, threads: int = 8, ahead: int = 16, rgb: bool = True, pin: bool = True):
        self.paths = [Path(p) for p in paths]
        self.threads, self.ahead, self.rgb = max(1, threads), max(1, ahead), rgb
        self.pin = pin and torch.cuda.is_available()
        self._free = {}

    def _buf(self, shape):
        pool = self._free.setdefault(shape, deque())
        if pool:
            return pool.popleft()
        t = torch.empty(shape, dtype=torch.uint8)
        return t.pin_memory() if self.pin else t

    def _decode(self, path):
        img = cv2.imread(str(path))
        if img is None:
            raise FileNotFoundError(f"Could not load image: {path}")
        return img

    def __len__(self):
        return len(self.paths)

    def __iter__(self) -> Iterator:
        live = deque()
        with ThreadPoolExecutor(self.threads) as ex:
            futs = deque()
            it = iter(enumerate(self.paths))
            for _ in range(self.ahead):
                nxt = next(it, None)
                if nxt is None:
                    break
                futs.append((nxt[0], ex.submit(self._decode, nxt[1])))
            while futs:
                i, fu = futs.popleft()
                img = fu.result()
                nxt = next(it, None)
                if nxt is not None:
                    futs.append((nxt[0], ex.submit(self._decode, nxt[1])))
                buf = self._buf(tuple(img.shape))
                if self.rgb:
                    cv2.cvtColor(img, cv2.COLOR_BGR2RGB, dst=buf.numpy())
                else:
                    buf.numpy()[...] = img
                live.append(buf)
                if len(live) > self.ahead + 2:
                    old = live.popleft()
                    self._free[tuple(old.shape)].append(old)
                yield i, buf
