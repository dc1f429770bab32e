#!/usr/bin/env python
"""One bench step (B pairs x 3 timesteps, 1080p, bf16, resident inputs) repeated a few times, for ncu:
    python tools/profile_step.py [steps] [pairs] [H W]
Prints the number of kernel launches per step so `ncu -s <skip> -c <count>` can isolate the last step.
Library options: VFI_OPT_<NAME>=<int> in the environment (e.g. VFI_OPT_T_FAST=0)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vfi_b200  # noqa
from vfi_b200 import _abi, ifnet_params
from vfi_b200.engine import Engine
from vfi_b200.synth import gen_clip

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
H, W = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (1080, 1920)
e = Engine.get(0)
e.load_weights(ifnet_params.random_init(1234))
for k, v in os.environ.items():                      # library options as VFI_OPT_<NAME>=<int>, e.g. VFI_OPT_T_FAST=0
    if k.startswith("VFI_OPT_"):
        e.set_option(k[8:].lower(), int(v))
clip = torch.from_numpy(gen_clip(B, H, W, seed=1, cut_every=0, fast=None)).cuda()
out = torch.empty((B, 3, H, W, 3), dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
l0 = e.launches
for k in range(steps):
    e.infer(clip[:-1], clip[1:], [0.25, 0.5, 0.75], 1.0, _abi.PRECISION_BF16, out=out)
torch.cuda.synchronize()
print("launches_per_step", (e.launches - l0) // steps, "checksum", int(out.sum()))
