#!/usr/bin/env python
"""Time the metric / routing-statistics kernels alone (bench.py's aux_kernels block): python tools/bench_aux.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import vfi_b200  # noqa: E402,F401
from vfi_b200.engine import Engine  # noqa: E402
from vfi_b200.synth import gen_clip  # noqa: E402

e = Engine.get(0)
clip = torch.from_numpy(gen_clip(24, 1080, 1920, seed=1, cut_every=0, fast=None)).cuda()
only = sys.argv[1:] or None
for k, v in bench.measure_aux(e, clip, 1080, 1920, bench.peaks()[0]).items():
    if only is None or any(o in k for o in only):
        print(f"{k:48s} {v['ms_per_frame'] * 1000:8.2f} us/frame  {v['achieved']:7.0f} GB/s  {v['frac']:.3f}")
