#!/usr/bin/env python
"""Time single conv layers through vfi_debug_conv with the library's per-class event timers (conversions excluded).
    python tools/bench_conv.py [ablate-mask ...]
The producer / MMA-warp masks (1, 2, 64, 128, 256, 512) act in every build; the EPILOGUE masks (4, 8, 16, 32, ...) exist only in the
single-group epilogue: build csrc with `make NVCC="nvcc -DV2_EPI_GROUPS=1"` for the phase matrix of DESIGN.md section 4b (the shipped
two-group epilogue carries no ablation branches)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vfi_b200  # noqa
from vfi_b200 import _abi, ifnet_params
from vfi_b200.engine import Engine
from vfi_b200.ifnet_layers import layer_info

e = Engine.get(0)
e.load_weights(ifnet_params.random_init(1234))
CASES = [(1, 8, 544, 960), (39, 12, 136, 240), (50, 12, 272, 480)]   # Head cnn1, block-3 ResConv, block-4 ResConv
args = sys.argv[1:]
if args and args[0] == "--layer":
    CASES = [c for c in CASES if c[0] == int(args[1])]
    args = args[2:]
masks = [int(a) for a in args] or [0]
NSCALE = float(os.environ.get("BC_NSCALE", "1"))   # scales the item count: slope = per-tile cost, intercept = launch cost
for layer, N, H, W in CASES:
    N = max(1, int(N * NSCALE))
    li = layer_info(layer)
    x = torch.randn((N, H, W, li["cin_pad"]), device="cuda")
    for m in masks:
        e.set_option("v2_ablate", m)
        for _ in range(2):
            e.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        e.profile(True)
        for _ in range(5):
            e.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        pr = e.profile_read()
        e.profile(False)
        ms, n = pr["conv_tcgen05"]
        print(f"layer {layer:2d} C={li['cin']:3d} {N}x{H}x{W} ablate={m:2d}: {ms / n * 1000:8.1f} us", flush=True)
    e.set_option("v2_ablate", 0)
