"""oracle/ifnet_ref.py: architecture checks, the golden fixture, and the wrapper
prologue/epilogue against the reference's own models/base.py outputs (tests/golden/base_helpers.npz)."""
import numpy as np
import torch
import torch.nn.functional as F

import vfi_b200  # noqa: F401
from oracle import ifnet_ref
from vfi_b200 import ifnet_params
from vfi_b200.synth import gen_clip


def test_param_count_and_strict_load(weights):
    net = ifnet_ref.make_ifnet()
    assert sum(p.numel() for p in net.parameters()) == ifnet_ref.PARAM_COUNT == ifnet_params.PARAM_COUNT
    net.load_state_dict(weights, strict=True)        # same keys/shapes as ifnet_params.param_spec()
    assert [k for k, _ in ifnet_params.param_spec()] == [k for k in weights]
    assert ifnet_params.flatten(weights).size == ifnet_params.PARAM_COUNT
    assert ifnet_params.flatten({"module." + k: v for k, v in weights.items()}).size == ifnet_params.PARAM_COUNT


def test_downscale_identity():
    """F.interpolate(1/s, bilinear, align_corners=False) == mean of the 2x2 centre pixels (App. B.1)."""
    x = torch.rand(1, 3, 64, 128)
    for s in (2, 4, 8, 16):
        o = s // 2 - 1
        want = (x[:, :, o::s, o::s] + x[:, :, o::s, o + 1::s] + x[:, :, o + 1::s, o::s] + x[:, :, o + 1::s, o + 1::s]) / 4
        got = F.interpolate(x, scale_factor=1.0 / s, mode="bilinear", align_corners=False)
        assert (got - want).abs().max() < 1e-6


def test_warp_is_clamped_pixel_offset_sampling():
    x = torch.arange(20, dtype=torch.float32).view(1, 1, 4, 5)
    fl = torch.zeros(1, 2, 4, 5)
    fl[:, 0] = 1.0                                   # sample at x+1: shifts left, right edge clamps
    w = ifnet_ref.warp(x, fl)
    assert torch.allclose(w[0, 0, :, :-1], x[0, 0, :, 1:], atol=1e-5) and torch.allclose(w[0, 0, :, -1], x[0, 0, :, -1], atol=1e-5)


def test_golden_fixture(golden, oracle_model):
    g = golden("ifnet_small.npz")
    n, H, W, seed = g["clip"].tolist()
    fr = gen_clip(n, H, W, seed=seed, cut_every=0, fast=None)
    u8, f32 = ifnet_ref.interpolate_ref(oracle_model, fr[0], fr[1], g["ts"].tolist(), return_float=True)
    assert np.abs(np.stack(f32) - g["out_f32"]).max() < 1e-5
    assert np.abs(np.stack(u8).astype(int) - g["out_u8"].astype(int)).max() <= 1


def test_prologue_epilogue_match_reference_base(golden):
    g = golden("base_helpers.npz")
    f = g["frame"]
    t = ifnet_ref.to_tensor(f)
    assert np.array_equal(t.numpy(), g["tensor"])
    tp = ifnet_ref.pad_reflect(t, 32)
    assert np.array_equal(tp.numpy(), g["padded"])
    assert np.array_equal(ifnet_ref.to_numpy_u8(tp[:, :, :f.shape[0], :f.shape[1]]), g["back"])
    assert np.array_equal(ifnet_ref.to_numpy_u8(t * 0 + 0.99999), g["trunc"]) and int(g["trunc"].max()) == 254
    assert ifnet_ref.pad_multiple_for(1.0) == 64 and ifnet_ref.pad_multiple_for(0.5) == 128


def test_model_surface(oracle_model):
    """The 4-method surface the reference loaders call (rife_wrapper.py:93-146)."""
    for name in ("load_model", "eval", "device", "inference"):
        assert callable(getattr(oracle_model, name))
    a = torch.rand(1, 3, 64, 64)
    out = oracle_model.inference(a, a.flip(3), timestep=0.5, scale=1.0)
    assert out.shape == (1, 3, 64, 64)
