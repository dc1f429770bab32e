#!/usr/bin/env python
"""Race hunt for the conv pipelines: the same step `reps` times at three shapes, every result compared bit for bit with the first.
    python tools/stress_determinism.py [reps]            the shipped kernels: must print RESULT ok
    python tools/stress_determinism.py [reps] nofence    with the epilogue's fence.proxy.async removed (v2_ablate bit 16): shows the
                                                         race that fence prevents (results change from run to run)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vfi_b200  # noqa
from vfi_b200 import _abi, ifnet_params
from vfi_b200.engine import Engine
from vfi_b200.synth import gen_clip

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
nofence = len(sys.argv) > 2 and sys.argv[2] == "nofence"
e = Engine.get(0)
e.load_weights(ifnet_params.random_init(1234))
bad = 0
for (B, H, W) in ((8, 1080, 1920), (30, 1080, 1920), (3, 1440, 2560)):
    clip = torch.from_numpy(gen_clip(B, H, W, seed=1, cut_every=0, fast=None)).cuda()
    ts = [0.25, 0.5, 0.75]
    ref = e.infer(clip[:-1], clip[1:], ts, 1.0, _abi.PRECISION_BF16).clone()   # reference: the shipped (fenced) kernels
    e.set_option("v2_ablate", 16 if nofence else 0)
    for k in range(reps if B == 8 else max(4, reps // 8)):
        out = e.infer(clip[:-1], clip[1:], ts, 1.0, _abi.PRECISION_BF16)
        if not torch.equal(out, ref):
            bad += 1
            print(f"MISMATCH {B}x{H}x{W} rep {k}: {(out != ref).sum().item()} bytes differ", flush=True)
    e.set_option("v2_ablate", 0)
    print(f"{B}x{H}x{W}: done", flush=True)
    del clip, ref, out
print("RESULT", "ok" if bad == 0 else f"{bad} mismatches")
sys.exit(0 if bad == 0 else 1)
