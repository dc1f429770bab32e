#!/usr/bin/env python
"""Flow-field statistics of the bench workload (random-init weights, synthetic clip): how coherent the gathers' taps are.
Measured: mean |flow| 0.57 px, max 3.5 px, mean |d flow / dx| 0.1 px/px."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import vfi_b200
from vfi_b200 import ifnet_params, _abi
from vfi_b200.engine import Engine
from vfi_b200.synth import gen_clip
e = Engine.get(0)
e.load_weights(ifnet_params.random_init(1234))
clip = gen_clip(2, 1080, 1920, seed=1)
c = torch.from_numpy(clip).cuda()
r = e.infer(c[:2], c[1:3], [0.25, 0.5, 0.75], 1.0, _abi.PRECISION_BF16, want_flow=True)
fl = r[-1] if isinstance(r, tuple) else None
print(type(r), [getattr(x, 'shape', None) for x in (r if isinstance(r, tuple) else [r])])
f = fl[0, 1].float()
print("flow mean abs", f.abs().mean().item(), "max", f.abs().max().item(), "std", f.std().item())
gx = (f[:, 1:] - f[:, :-1]).abs().mean().item(); gy = (f[1:] - f[:-1]).abs().mean().item()
print("mean |d flow/dx|", gx, "mean |d flow/dy|", gy)
