import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vfi_b200  # noqa
from vfi_b200 import _abi, ifnet_params
from vfi_b200.engine import Engine
from vfi_b200.ifnet_layers import layer_info
e = Engine.get(0)
e.load_weights(ifnet_params.random_init(1234))
CASES = [(1, 9, 544, 960), (3, 9, 544, 960), (17, 24, 34, 60), (28, 24, 68, 120), (39, 24, 136, 240), (50, 24, 272, 480), (58, 24, 272, 480)]
for layer, N, H, W in CASES:
    li = layer_info(layer)
    x = torch.randn((N, H, W, li["cin_pad"]), device="cuda")
    for tma in (0, 1, 0, 1):
        e.set_option("v2_tma", tma)
        for _ in range(2):
            e.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        e.profile(True)
        for _ in range(5):
            e.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        pr = e.profile_read()
        e.profile(False)
        ms, n = pr["conv_tcgen05"]
        print(f"layer {layer:2d} C={li['cin']:3d} {N}x{H}x{W} tma={tma}: {ms / n * 1000:8.1f} us", flush=True)
e.set_option("v2_tma", 1)
