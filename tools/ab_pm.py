#!/usr/bin/env python
"""A/B of conv_v2 tile layouts for the c = 64 ResConv (layer 39): chunk-major 16-byte TMA rows vs pixel-major
128-byte-swizzled rows, with the ablation masks of tools/bench_conv.py (2: no MMAs, 4: no epilogue math/stores)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vfi_b200  # noqa
from vfi_b200 import _abi, ifnet_params
from vfi_b200.engine import Engine
from vfi_b200.ifnet_layers import layer_info
e = Engine.get(0)
e.load_weights(ifnet_params.random_init(1234))
layer, N, H, W = 39, 24, 136, 240
li = layer_info(layer)
x = torch.randn((N, H, W, li["cin_pad"]), device="cuda")
for pm in (0, 1, 0, 1):
    e.set_option("v2_pm", pm)
    for m in (0, 2, 4, 6):
        e.set_option("v2_ablate", m)
        for _ in range(2):
            e.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        e.profile(True)
        for _ in range(5):
            e.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        pr = e.profile_read()
        e.profile(False)
        ms, n = pr["conv_tcgen05"]
        print(f"pm={pm} ablate={m}: {ms / n * 1000:8.1f} us", flush=True)
e.set_option("v2_ablate", 0)
e.set_option("v2_pm", 1)
