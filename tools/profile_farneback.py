#!/usr/bin/env python
"""Farneback (reference parameters) on 8 1080p pairs, repeated, for an ncu launch list / CUDA-event timing:
    python tools/profile_farneback.py [reps] [pairs]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vfi_b200  # noqa
from vfi_b200.engine import Engine
from vfi_b200.synth import gen_clip

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8
e = Engine.get(0)
clip = torch.from_numpy(gen_clip(n, 1080, 1920, seed=2, cut_every=0, fast=None)).cuda()
torch.cuda.synchronize()
l0 = e.launches
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for k in range(reps):
    if k == reps - 1:
        e0.record()
    st = e.farneback(clip[:-1], clip[1:], want_flow=False, want_stats=True)
e1.record()
torch.cuda.synchronize()
print("launches_per_call", (e.launches - l0) // reps, "ms_per_pair", e0.elapsed_time(e1) / n, "stats0", st[0].tolist())
