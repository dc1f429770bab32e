#!/usr/bin/env python
"""GPU bring-up diagnostics: per-layer conv parity (direct fp32 and tcgen05 bf16) against torch,
whole-net parity against the CPU oracle, reductions against the oracle.  Prints one line per check;
never raises, so a single gpurun call shows everything that is wrong."""
import json
import os
import sys
import time
import traceback

import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vfi_b200  # noqa
from vfi_b200 import _abi, ifnet_params
from vfi_b200.engine import Engine, HudState
from vfi_b200.ifnet_layers import layer_info
from vfi_b200.synth import gen_clip, noisy_copy

OUT = {}
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
PRECS = os.environ.get("VFI_DIAG_PREC", "fp32,bf16").split(",")


def rec(name, **kw):
    OUT[name] = kw
    print(name, json.dumps(kw), flush=True)


def bf16r(t):
    return t.to(torch.bfloat16).to(torch.float32)


def torch_layer(sd, layer, x_nhwc, res_nhwc, slope, round_bf16):
    li = layer_info(layer)
    w = sd[li["key"] + ".weight"].cuda()
    b = sd[li["key"] + ".bias"].cuda()
    x = x_nhwc[..., :li["cin"]].permute(0, 3, 1, 2).contiguous()
    if round_bf16:
        w, x = bf16r(w), bf16r(x)
    if li["kind"] == "conv":
        y = F.conv2d(x, w, b, stride=li["stride"], padding=1)
        if li["beta"]:
            r = res_nhwc.permute(0, 3, 1, 2)
            if round_bf16:
                r = bf16r(r)
            y = y * sd[li["beta_key"]].cuda() + r
        if li["act"]:
            y = F.leaky_relu(y, slope)
        return y.permute(0, 2, 3, 1).contiguous()
    y = F.conv_transpose2d(x, w, b, stride=2, padding=1)
    if layer == 3:
        return y.permute(0, 2, 3, 1).contiguous()
    y = F.pixel_shuffle(y, 2)
    return y.permute(0, 2, 3, 1).contiguous()  # 13 channels


def conv_checks(e, sd, layers, Hin=24, Win=40, N=2):
    g = torch.Generator(device="cpu").manual_seed(0)
    for layer in layers:
        li = layer_info(layer)
        for prec, name in ((_abi.PRECISION_FP32, "fp32"), (_abi.PRECISION_BF16, "bf16")):
            if name not in PRECS:
                continue
            try:
                x = torch.randn((N, Hin, Win, li["cin_pad"]), generator=g).cuda()
                x[..., li["cin"]:] = 0
                res = None
                if li["beta"]:
                    res = torch.randn((N, Hin, Win, li["cout"]), generator=g).cuda()
                y = e.debug_conv(layer, prec, x, res)
                torch.cuda.synchronize()
                ref = torch_layer(sd, layer, x, res, 0.2, prec == _abi.PRECISION_BF16)
                yy = y[..., :ref.shape[-1]]
                err = (yy - ref).abs().max().item()
                scale = ref.abs().max().item()
                rec(f"conv_L{layer}_{name}", key=li["key"], cin=li["cin"], cout=li["cout"], stride=li["stride"],
                    max_err=err, ref_max=scale, ok=bool(err <= (2e-2 if prec else 1e-4) * max(scale, 1.0)))
            except Exception as ex:
                rec(f"conv_L{layer}_{name}", error=repr(ex)[:300])
                traceback.print_exc()


def net_checks(e, sd, H, W, ts, scale=1.0):
    from oracle import ifnet_ref
    fr = gen_clip(1, H, W, seed=0, cut_every=0, fast=None)
    m = ifnet_ref.Model()
    m.flownet.load_state_dict(sd, strict=True)
    t0 = time.time()
    ref_u8, ref_f = ifnet_ref.interpolate_ref(m, fr[0], fr[1], ts, scale=scale, return_float=True)
    cpu_s = time.time() - t0
    for prec, name, tol in ((_abi.PRECISION_FP32, "fp32", 1e-4), (_abi.PRECISION_BF16, "bf16", 2e-2)):
        if name not in PRECS:
            continue
        try:
            out, of, fl = e.infer(fr[0][None], fr[1][None], ts, scale, prec, want_float=True, want_flow=True)
            torch.cuda.synchronize()
            of = of[0].cpu().numpy()
            errs = [float(np.abs(of[i] - ref_f[i]).max()) for i in range(len(ts))]
            mse = [float(((of[i] - ref_f[i]) ** 2).mean()) for i in range(len(ts))]
            psnr = [10 * np.log10(1.0 / max(v, 1e-20)) for v in mse]
            du8 = int(np.abs(out[0].cpu().numpy().astype(int) - np.stack(ref_u8).astype(int)).max())
            rec(f"net_{H}x{W}_s{scale}_{name}", max_abs=errs, psnr_vs_ref=psnr, u8_maxdiff=du8,
                flow_absmax=float(fl.abs().max()), ok=bool(max(errs) <= tol), oracle_cpu_s=cpu_s)
        except Exception as ex:
            rec(f"net_{H}x{W}_s{scale}_{name}", error=repr(ex)[:300])
            traceback.print_exc()


def reduce_checks(e):
    from oracle import metrics_ref, router_ref
    fr = gen_clip(7, 360, 640, seed=3, cut_every=4, fast=None)
    pred = noisy_copy(fr, seed=5)
    try:
        sse, ps = e.psnr_sse(pred, fr)
        ref = [int(((pred[i].astype(np.int64) - fr[i].astype(np.int64)) ** 2).sum()) for i in range(len(fr))]
        rec("psnr", ok=bool(sse.tolist() == ref), psnr0=float(ps[0]), ref0=metrics_ref.psnr_simple(pred[0], fr[0]))
    except Exception as ex:
        rec("psnr", error=repr(ex)[:300]); traceback.print_exc()
    try:
        s1 = e.ssim(pred, fr, 1).cpu().numpy()
        r1 = np.array([metrics_ref.ssim_simple(pred[i], fr[i]) for i in range(len(fr))])
        s4 = e.ssim(fr[:-1], fr[1:], 4).cpu().numpy()
        r4 = np.array([router_ref.scene_ssim(fr[i], fr[i + 1]) for i in range(len(fr) - 1)])
        rec("ssim", err_full=float(np.abs(s1 - r1).max()), err_quarter=float(np.abs(s4 - r4).max()),
            quarter=r4.tolist(), ok=bool(np.abs(s1 - r1).max() < 1e-9 and np.abs(s4 - r4).max() < 1e-9))
    except Exception as ex:
        rec("ssim", error=repr(ex)[:300]); traceback.print_exc()
    try:
        import cv2
        g0 = cv2.cvtColor(fr[0], cv2.COLOR_RGB2GRAY); g1 = cv2.cvtColor(fr[1], cv2.COLOR_RGB2GRAY)
        flow = router_ref.farneback(g0, g1)
        st = e.flow_stats(torch.from_numpy(flow)[None])[0].tolist()
        rm, rx, rs, _ = router_ref.flow_stats(flow)
        rec("flow_stats", got=st, ref=[rm, rx, rs],
            ok=bool(abs(st[0] - rm) <= 1e-5 * abs(rm) + 1e-9 and st[1] == rx and abs(st[2] - rs) <= 1e-5 * abs(rs) + 1e-9))
    except Exception as ex:
        rec("flow_stats", error=repr(ex)[:300]); traceback.print_exc()
    try:
        lv = e.laplacian_var(fr).cpu().numpy()
        rl = np.array([router_ref.laplacian_var(f) for f in fr])
        rec("laplacian", err=float(np.abs(lv - rl).max()), ok=bool(np.allclose(lv, rl, rtol=1e-12, atol=1e-9)))
    except Exception as ex:
        rec("laplacian", error=repr(ex)[:300]); traceback.print_exc()
    try:
        r = router_ref.RouterRef()
        hs = HudState(e)
        okm, covs = True, []
        clip = gen_clip(8, 360, 640, seed=4, cut_every=0, fast=None)
        for f in clip:
            m_ref, c_ref = r.detect_hud(f)
            m, c = hs.update(f)
            m = m.cpu().numpy().astype(bool)
            same = bool((m == m_ref).all()) and float(c[0]) == c_ref
            okm &= same
            covs.append([float(c[0]), c_ref])
        rec("hud", ok=okm, cov=covs)
    except Exception as ex:
        rec("hud", error=repr(ex)[:300]); traceback.print_exc()


def main():
    which = sys.argv[1:] or ["conv", "net", "reduce"]
    e = Engine.get(0)
    sd = ifnet_params.random_init(1234)
    e.load_weights(sd)
    rec("load", ok=True, gpu=torch.cuda.get_device_name(0))
    if "conv" in which:
        conv_checks(e, sd, [1, 3, 4, 5, 6, 14, 15, 16, 17, 25, 26, 27, 28, 36, 37, 38, 39, 47, 48, 49, 50, 58])
    if "reduce" in which:
        reduce_checks(e)
    if "net" in which:
        net_checks(e, sd, 128, 192, [0.5])
        net_checks(e, sd, 270, 480, [0.25, 0.5, 0.75])
    if "net1080" in which:
        net_checks(e, sd, 1080, 1920, [0.5])
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(OUT, open("gpurun_out/diag_%s.json" % "_".join(PRECS), "w"), indent=1)
    bad = [k for k, v in OUT.items() if not v.get("ok", False)]
    print("FAILED:", bad)


if __name__ == "__main__":
    main()
