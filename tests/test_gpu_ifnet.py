"""Parity of the CUDA IFNet pass (through the C ABI) with the CPU oracle.
Tolerances are BASELINE.json north_star's: fp32 mode max-abs 1e-4 on [0,1] frames; bf16 mode max-abs 2e-2 and
PSNR-vs-oracle >= 45 dB.  At 1080p (too slow for the oracle inside a test) size-independent properties are used."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import vfi_b200  # noqa: F401
from oracle import ifnet_ref
from vfi_b200 import _abi
from vfi_b200.ifnet_layers import layer_info
from vfi_b200.synth import gen_clip

pytestmark = pytest.mark.gpu
TOL = {_abi.PRECISION_FP32: 1e-4, _abi.PRECISION_BF16: 2e-2}
TS = [0.25, 0.5, 0.75]


def _psnr(a, b):
    return 10 * np.log10(1.0 / max(float(((a - b) ** 2).mean()), 1e-20))


def _check(engine, oracle_model, fr0, fr1, ts, scale, prec):
    ref_u8, ref_f = ifnet_ref.interpolate_ref(oracle_model, fr0, fr1, ts, scale=scale, return_float=True)
    out, of = engine.infer(fr0[None], fr1[None], ts, scale, prec, want_float=True)
    of, out = of[0].cpu().numpy(), out[0].cpu().numpy()
    for i in range(len(ts)):
        err = float(np.abs(of[i] - ref_f[i]).max())
        assert err <= TOL[prec], f"t={ts[i]} max-abs {err} > {TOL[prec]}"
        if prec == _abi.PRECISION_BF16:
            assert _psnr(of[i], ref_f[i]) >= 45.0
        # uint8 store = truncation of the float result (base.py:201); vs the oracle's u8 at most 1 + tol*255 levels
        assert np.array_equal(out[i], (of[i] * 255).clip(0, 255).astype(np.uint8))
        assert np.abs(out[i].astype(int) - ref_u8[i].astype(int)).max() <= int(TOL[prec] * 255) + 1


@pytest.mark.parametrize("prec", [_abi.PRECISION_FP32, _abi.PRECISION_BF16])
def test_golden_fixture(engine, golden, prec):
    g = golden("ifnet_small.npz")
    n, H, W, seed = g["clip"].tolist()
    fr = gen_clip(n, H, W, seed=seed, cut_every=0, fast=None)
    out, of = engine.infer(fr[0][None], fr[1][None], g["ts"].tolist(), 1.0, prec, want_float=True)
    assert float(np.abs(of[0].cpu().numpy() - g["out_f32"]).max()) <= TOL[prec]


@pytest.mark.parametrize("prec", [_abi.PRECISION_FP32, _abi.PRECISION_BF16])
@pytest.mark.parametrize("H,W", [(128, 192), (270, 480), (100, 150)])     # aligned, 1080p/4, ragged (reflect pad both ways)
def test_parity_scale1(engine, oracle_model, prec, H, W):
    fr = gen_clip(1, H, W, seed=0, cut_every=0, fast=None)
    _check(engine, oracle_model, fr[0], fr[1], TS, 1.0, prec)


@pytest.mark.parametrize("prec", [_abi.PRECISION_FP32, _abi.PRECISION_BF16])
@pytest.mark.parametrize("scale,H,W", [(0.5, 256, 384), (0.5, 200, 300), (0.25, 256, 512)])
def test_parity_uhd_scales(engine, oracle_model, prec, scale, H, W):
    fr = gen_clip(1, H, W, seed=7, cut_every=0, fast=None)
    _check(engine, oracle_model, fr[0], fr[1], [0.5], scale, prec)


@pytest.mark.parametrize("opts", [{"v2_tma": 0}, {"fuse_gather": 0}, {"conv_v2": 0}, {"v2_res_global": 1}, {"l2_chunk_mb": 1}, {"v2_pm": 0},
                                  {"fuse_gather3": 1}, {"gc_persistent": 0}, {"t_fast": 0}],
                         ids=lambda o: ",".join(f"{k}={v}" for k, v in o.items()))
def test_parity_of_every_kernel_variant(engine, oracle_model, opts):
    """vfi_set_option only chooses between kernels that compute the same thing: each non-default path (cp.async
    producers, unfused finest gather + tiled blend, TMA-per-tap convs, residual from global memory, depth-first
    ResConv chain, chunk-major c = 64 tiles, block 3's gather fused with its conv0[0], the one-tile-per-CTA form of the
    fused finest gather, the item-major tile order of the gathers and the blend) meets the bf16 tolerance against the oracle on its own."""
    defaults = {"v2_tma": 1, "fuse_gather": 1, "conv_v2": 1, "v2_res_global": 0, "l2_chunk_mb": 0, "v2_pm": 1, "fuse_gather3": 0,
                "gc_persistent": 1, "t_fast": 1}
    fr = gen_clip(2, 160, 224, seed=11, cut_every=0, fast=None)
    try:
        for k, v in opts.items():
            engine.set_option(k, v)
        _check(engine, oracle_model, fr[0], fr[1], TS, 1.0, _abi.PRECISION_BF16)
        two = engine.infer(fr[:2], fr[1:3], TS, 1.0, _abi.PRECISION_BF16).cpu().numpy()       # batch of 2: item groups
        one = engine.infer(fr[1][None], fr[2][None], TS, 1.0, _abi.PRECISION_BF16).cpu().numpy()
        assert np.array_equal(two[1], one[0])
    finally:
        for k, v in defaults.items():
            engine.set_option(k, v)


@pytest.mark.parametrize("prec", [_abi.PRECISION_FP32, _abi.PRECISION_BF16])
@pytest.mark.parametrize("other", [{}, {"gc_persistent": 0}, {"fuse_gather": 0}], ids=["default", "one-tile-gather", "unfused-gather"])
def test_tile_order_changes_no_bit(engine, prec, other):
    """"t_fast" only permutes the order in which the gathers and the blend visit their (tile, item) pairs (timestep-fastest, so a
    pair's images are read once for all T items): frames and flow are identical to the item-major order, bit for bit."""
    fr = gen_clip(3, 200, 300, seed=13, cut_every=0, fast=None)
    try:
        for k, v in other.items():
            engine.set_option(k, v)
        got = [t.clone() for t in engine.infer(fr[:3], fr[1:4], [0.25, 0.5, 0.75, 0.9], 1.0, prec, want_flow=True)]
        for order in (0, 2):                                   # item-major; (timestep, pair, x, y) in the fused finest gather
            engine.set_option("t_fast", order)
            want = engine.infer(fr[:3], fr[1:4], [0.25, 0.5, 0.75, 0.9], 1.0, prec, want_flow=True)
            assert torch.equal(got[0], want[0]) and torch.equal(got[1], want[1]), order
    finally:
        for k, v in {"t_fast": 1, "gc_persistent": 1, "fuse_gather": 1}.items():
            engine.set_option(k, v)


def test_timestep_edge_values(engine, oracle_model):
    fr = gen_clip(1, 128, 192, seed=3, cut_every=0, fast=None)
    _check(engine, oracle_model, fr[0], fr[1], [0.0, 1.0, 0.125], 1.0, _abi.PRECISION_FP32)


@pytest.mark.parametrize("prec", [_abi.PRECISION_FP32, _abi.PRECISION_BF16])
def test_batch_equals_singles_bit_exact(engine, prec):
    fr = gen_clip(3, 192, 256, seed=5, cut_every=0, fast=None)
    batch = engine.infer(fr[:3], fr[1:4], TS, 1.0, prec).cpu().numpy()
    for b in range(3):
        one = engine.infer(fr[b][None], fr[b + 1][None], TS, 1.0, prec).cpu().numpy()
        assert np.array_equal(batch[b], one[0])
        for i, t in enumerate(TS):      # and each timestep on its own (Head sharing does not change results)
            assert np.array_equal(engine.infer(fr[b][None], fr[b + 1][None], [t], 1.0, prec).cpu().numpy()[0, 0], one[0, i])


def test_hud_mask_composite(engine):
    fr = gen_clip(1, 128, 192, seed=2, cut_every=0, fast=None)
    m = np.zeros((1, 128, 192), np.uint8)
    m[0, :20, :50] = 1
    plain = engine.infer(fr[0][None], fr[1][None], TS, 1.0, _abi.PRECISION_BF16).cpu().numpy()[0]
    comp = engine.infer(fr[0][None], fr[1][None], TS, 1.0, _abi.PRECISION_BF16, hud_mask=m).cpu().numpy()[0]
    mb = m[0].astype(bool)
    for i, t in enumerate(TS):          # adaptive_pipeline.py:399-409
        assert np.array_equal(comp[i][mb], (fr[0] if t < 0.5 else fr[1])[mb])
        assert np.array_equal(comp[i][~mb], plain[i][~mb])


def test_argument_errors(engine):
    z = np.zeros((1, 64, 64, 3), np.uint8)
    with pytest.raises(_abi.VFIError, match="scale"):
        engine.infer(z, z, [0.5], 0.3, _abi.PRECISION_BF16)
    with pytest.raises(_abi.VFIError):
        engine.infer(z[:, :1], z[:, :1], [0.5], 1.0, _abi.PRECISION_BF16)       # too small to reflect-pad


LAYERS = [1, 3, 4, 5, 6, 13, 14, 15, 16, 17, 25, 26, 27, 28, 36, 37, 38, 39, 47, 48, 49, 50, 57, 58]


def _torch_layer(sd, layer, x, res, round_bf16):
    li = layer_info(layer)
    w, b = sd[li["key"] + ".weight"].cuda(), sd[li["key"] + ".bias"].cuda()
    x = x[..., :li["cin"]].permute(0, 3, 1, 2).contiguous()
    r16 = (lambda t: t.to(torch.bfloat16).float()) if round_bf16 else (lambda t: t)
    w, x = r16(w), r16(x)
    if li["kind"] == "conv":
        y = F.conv2d(x, w, b, stride=li["stride"], padding=1)
        if li["beta"]:
            y = y * sd[li["beta_key"]].cuda() + (r16(res.permute(0, 3, 1, 2)) if res is not None else x)
        if li["act"]:
            y = F.leaky_relu(y, 0.2)
        return y.permute(0, 2, 3, 1)
    y = F.conv_transpose2d(x, w, b, stride=2, padding=1)
    return (y if layer == 3 else F.pixel_shuffle(y, 2)).permute(0, 2, 3, 1)


@pytest.mark.parametrize("prec", [_abi.PRECISION_FP32, _abi.PRECISION_BF16])
@pytest.mark.parametrize("layer", LAYERS)
def test_single_layer_vs_torch(engine, weights, layer, prec):
    """Every conv kernel shape of the net against torch fp32 conv2d / conv_transpose2d on the same
    (bf16-rounded, for the tcgen05 path) operands; odd spatial sizes exercise partial tiles."""
    torch.backends.cudnn.allow_tf32 = False
    li = layer_info(layer)
    g = torch.Generator().manual_seed(layer)
    Hin, Win = (26, 44) if li["stride"] == 2 and li["kind"] == "conv" else (23, 41)
    x = torch.randn((2, Hin, Win, li["cin_pad"]), generator=g).cuda()
    x[..., li["cin"]:] = 0
    res = torch.randn((2, Hin, Win, li["cout"]), generator=g).cuda() if li["beta"] else None
    y = engine.debug_conv(layer, prec, x, res)
    ref = _torch_layer(weights, layer, x, res, prec == _abi.PRECISION_BF16)
    nc = 5 if (layer == 58 and prec == _abi.PRECISION_BF16) else ref.shape[-1]    # final lastconv: only flow + mask are computed
    ref = ref[..., :nc]
    err = (y[..., :nc] - ref).abs().max().item()
    tol = (1e-4 if prec == _abi.PRECISION_FP32 else 1.5e-2) * max(1.0, ref.abs().max().item())
    assert err <= tol, f"layer {layer} ({li['key']}): {err} > {tol}"


@pytest.mark.parametrize("layer", [1, 2, 17, 24, 28, 35, 39, 46, 50, 57])
@pytest.mark.parametrize("Hin,Win", [(16, 8), (23, 41), (50, 100), (136, 240)])
def test_conv_v2_halo_tile_kernel(engine, weights, layer, Hin, Win):
    """csrc/conv_v2.cu (resident weights, one halo tile per 16 x 8KW outputs, shifted no-swizzle descriptors) against
    torch and against the TMA-per-tap kernel on the same bf16 operands; partial tiles in both directions."""
    torch.backends.cudnn.allow_tf32 = False
    li = layer_info(layer)
    g = torch.Generator().manual_seed(100 + layer)
    x = torch.randn((3, Hin, Win, li["cin_pad"]), generator=g).cuda()
    try:
        engine.set_option("conv_v2", 1)
        y2 = engine.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        engine.set_option("conv_v2", 0)
        y1 = engine.debug_conv(layer, _abi.PRECISION_BF16, x, None)
    finally:
        engine.set_option("conv_v2", 1)
    ref = _torch_layer(weights, layer, x, None, True)
    scale = max(1.0, ref.abs().max().item())
    assert (y2 - ref).abs().max().item() <= 1.5e-2 * scale
    assert (y2 - y1).abs().max().item() <= 1.0e-2 * scale          # same operands, different accumulation order + bf16 store
    if li["cin"] == 64:     # the chunk-major (default: pixel-major SWIZZLE_128B, tap-shifted descriptors) and cp.async-fed variants:
        for opt in ("v2_pm", "v2_tma"):   # same MMAs, same bits
            try:
                engine.set_option(opt, 0)
                y3 = engine.debug_conv(layer, _abi.PRECISION_BF16, x, None)
            finally:
                engine.set_option(opt, 1)   # both default to 1
            assert torch.equal(y3, y2), opt


@pytest.mark.parametrize("layer", [15, 26, 27, 37, 38, 48, 49])
@pytest.mark.parametrize("Hin,Win", [(26, 44), (64, 96), (136, 240)])
def test_conv_s2_plane_window_kernel(engine, weights, layer, Hin, Win):
    """csrc/conv_s2.cu (stride-2 convs: resident weights, one 17 x (8KW+1) window per phase plane per tile) against
    torch and against the TMA-per-tap kernel on the same bf16 operands; partial tiles in both directions."""
    torch.backends.cudnn.allow_tf32 = False
    li = layer_info(layer)
    g = torch.Generator().manual_seed(200 + layer)
    x = torch.randn((2, Hin, Win, li["cin_pad"]), generator=g).cuda()
    x[..., li["cin"]:] = 0
    try:
        engine.set_option("conv_v2", 1)
        y2 = engine.debug_conv(layer, _abi.PRECISION_BF16, x, None)
        engine.set_option("conv_v2", 0)
        y1 = engine.debug_conv(layer, _abi.PRECISION_BF16, x, None)
    finally:
        engine.set_option("conv_v2", 1)
    ref = _torch_layer(weights, layer, x, None, True)
    scale = max(1.0, ref.abs().max().item())
    assert (y2 - ref).abs().max().item() <= 1.5e-2 * scale
    assert (y2 - y1).abs().max().item() <= 1.0e-2 * scale
