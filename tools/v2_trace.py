#!/usr/bin/env python
"""Pipeline time-line of CTA 0 of one conv_v2 launch (vfi_set_option "v2_trace"):
    python tools/v2_trace.py run     # on the GPU box: traces layers 50 (C=32) and 39 (C=64)
    python tools/v2_trace.py show gpurun_out/v2_trace_c32.bin"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def show(path, ntiles=24):
    t = np.fromfile(path, dtype=np.uint64).reshape(3, 64, 8).astype(np.int64)
    t0 = t[t > 0].min()
    r = np.where(t > 0, t - t0, -1)
    print(path)
    print("tile | prod: want  got  issued  published | mma: start accfree landed issued | epi: begin visible res-read accdone ld0 ld1 released stored  (SM cycles)")
    for k in range(ntiles):
        if r[0, k, 0] < 0 and r[1, k, 0] < 0:
            break
        print(f"{k:4d} | " + " ".join(f"{v:7d}" for v in r[0, k, :4]) + " | " + " ".join(f"{v:7d}" for v in r[1, k, :4]) +
              " | " + " ".join(f"{r[2, k, e]:7d}" for e in (3, 4, 0, 1, 5, 6, 2, 7)))


if sys.argv[1] == "show":
    show(sys.argv[2])
else:
    import torch
    import vfi_b200  # noqa
    from vfi_b200 import _abi, ifnet_params
    from vfi_b200.engine import Engine
    from vfi_b200.ifnet_layers import layer_info
    e = Engine.get(0)
    e.load_weights(ifnet_params.random_init(1234))
    NN = int(os.environ.get("V2_TRACE_N", "12"))
    for layer, N, H, W in [(50, NN, 272, 480), (39, NN, 136, 240)]:
        li = layer_info(layer)
        x = torch.randn((N, H, W, li["cin_pad"]), device="cuda")
        for _ in range(2):
            e.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        e.set_option("v2_ablate", int(os.environ.get("V2_TRACE_ABLATE", "0")))
        e.set_option("v2_trace", 1)
        e.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        e.set_option("v2_trace", 0)
        e.set_option("v2_ablate", 0)
        show(f"gpurun_out/v2_trace_c{li['cin']}.bin")
