"""IFNet v4.25 parameter inventory: names, shapes, canonical flat order (include/vfi_b200.h,
vfi_ifnet_load_weights), random initialisation and flownet.pkl loading.

Key names are the v4.25 state_dict's; the tests load a dict produced here
into the oracle module with strict=True, which is the architecture check.
"""
import math
import os

import numpy as np
import torch

BLOCK_C = (192, 128, 96, 64, 32)
PARAM_COUNT = 5_664_776


def param_spec():
    """Ordered [(key, shape)] in the flat order the C ABI expects."""
    spec = [("encode.cnn0.weight", (16, 3, 3, 3)), ("encode.cnn0.bias", (16,)),
            ("encode.cnn1.weight", (16, 16, 3, 3)), ("encode.cnn1.bias", (16,)),
            ("encode.cnn2.weight", (16, 16, 3, 3)), ("encode.cnn2.bias", (16,)),
            ("encode.cnn3.weight", (16, 4, 4, 4)), ("encode.cnn3.bias", (4,))]
    for b, c in enumerate(BLOCK_C):
        cin = 15 if b == 0 else 28
        p = f"block{b}."
        spec += [(p + "conv0.0.0.weight", (c // 2, cin, 3, 3)), (p + "conv0.0.0.bias", (c // 2,)),
                 (p + "conv0.1.0.weight", (c, c // 2, 3, 3)), (p + "conv0.1.0.bias", (c,))]
        for j in range(8):
            spec += [(p + f"convblock.{j}.conv.weight", (c, c, 3, 3)), (p + f"convblock.{j}.conv.bias", (c,)),
                     (p + f"convblock.{j}.beta", (1, c, 1, 1))]
        spec += [(p + "lastconv.0.weight", (c, 52, 4, 4)), (p + "lastconv.0.bias", (52,))]
    return spec


def random_init(seed=1234):
    """PyTorch-default-style init (kaiming-uniform a=sqrt(5): U(-1/sqrt(fan_in), 1/sqrt(fan_in)), beta = 1)
    from a seeded numpy generator; a deterministic weight choice for the synthetic configs."""
    rng = np.random.default_rng(seed)
    sd = {}
    fan = 1
    for key, shape in param_spec():
        if key.endswith("beta"):
            sd[key] = torch.ones(shape, dtype=torch.float32)
            continue
        if key.endswith("weight"):
            fan = shape[1] * shape[2] * shape[3]
        bound = 1.0 / math.sqrt(fan)
        sd[key] = torch.from_numpy(rng.uniform(-bound, bound, shape).astype(np.float32))
    return sd


def flatten(sd):
    """state_dict -> float32 vector in canonical order (accepts a leading 'module.' on keys)."""
    sd = {(k[7:] if k.startswith("module.") else k): v for k, v in sd.items()}
    parts = []
    for key, shape in param_spec():
        if key not in sd:
            raise KeyError(f"IFNet state_dict is missing {key}")
        t = sd[key].detach().to(torch.float32).cpu().contiguous()
        if tuple(t.shape) != tuple(shape):
            raise ValueError(f"{key}: expected shape {shape}, got {tuple(t.shape)}")
        parts.append(t.reshape(-1).numpy())
    flat = np.ascontiguousarray(np.concatenate(parts), dtype=np.float32)
    assert flat.size == PARAM_COUNT
    return flat


def load_pkl(path):
    """Model.load_model(path, -1) semantics: <path>/flownet.pkl, strip 'module.' (Appendix A.1)."""
    f = os.path.join(str(path), "flownet.pkl")
    sd = torch.load(f, map_location="cpu")
    return {(k[7:] if k.startswith("module.") else k): v for k, v in sd.items()}
