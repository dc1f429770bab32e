#!/usr/bin/env python
"""One pair per call (the reference's call shape, scripts/run_benchmarks.py:142-149): P50 / P99 of RIFEModel.interpolate and
AdaptivePipeline.process_pair with host numpy frames in and out.  VFI_B200_NO_GRAPH=1 disables the CUDA-graph replay."""
import contextlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vfi_b200  # noqa
from vfi_b200.models.sota.rife_wrapper import RIFEModel
from vfi_b200.synth import gen_clip

m = RIFEModel(allow_random_weights=True)
with contextlib.redirect_stdout(sys.stderr):
    m.load()
for H, W in ((360, 640), (720, 1280), (1080, 1920)):
  fr = gen_clip(4, H, W, seed=1, cut_every=0, fast=None)
  for T in (1, 3):
    ts = []
    for k in range(60):
        t0 = time.perf_counter()
        out = m.interpolate(fr[k % 4], fr[k % 4 + 1], T)
        ts.append((time.perf_counter() - t0) * 1000)
    ts = np.array(ts[10:])
    print(f"{H}x{W} interpolate T={T} graph={m.use_graph}: p50 {np.percentile(ts, 50):.3f} ms  p99 {np.percentile(ts, 99):.3f} ms  checksum {int(out[0].sum())}")
