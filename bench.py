#!/usr/bin/env python
"""bench.py -- interpolated 1080p frames/s, RIFE v4.25 4x (t = 0.25/0.5/0.75), bf16, synthetic clip.

    python bench.py [--gpus N] [--steps K] [--warmup W]            (N > 1: launched under torchrun)
    python bench.py --impl reference ...                           (CPU oracle on the host cores)

One "step" = one pass of the hot path over one batch of `--pairs` frame pairs x 3 timesteps
(BASELINE.json configs[1]; the 300-pair clip = 10 steps of 30 pairs).  `value` is device-timed with the clip
resident in HBM; `e2e` goes through RIFEModel.clip_streamer with pinned HOST buffers (H2D + D2H
inside the timed region).  `other_configs` carries BASELINE.json configs 1, 3, 4 and 5 (each sharded over
the ranks like the headline, with a parity block checked against the committed reference-generated
golden / exact integer sums) and `cpu_baseline` the three host-side baselines.  Prints ONE JSON line
on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W = 1080, 1920
TIMESTEPS = [0.25, 0.5, 0.75]
METRIC = "interpolated_1080p_frames_per_sec_rife_4x"
UNIT = "frames/s"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


def ncu_traffic(pattern):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes, average over the step's launches whose kernel name
    matches `pattern`) from the newest committed ncu launch list of this workload (profiles/r0N_launches_final.txt:
    one step of the default batch, 30 pairs x 3 timesteps of 1080p frames); None if no such file."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r0*_launches_final.txt")))
    if not files:
        return None
    tot, n = 0.0, 0
    for ln in open(files[-1]):
        if ln.startswith("#") or not re.search(pattern, ln):
            continue
        m = re.search(r"rd\s+([\d.]+) MB\s+wr\s+([\d.]+) MB", ln)
        if m:
            tot += (float(m.group(1)) + float(m.group(2))) * 1e6
            n += 1
    return tot / n if n else None


def tc_flops_per_batch(B, T, Hp, Wp, n_images, fused_conv0a):
    """Algorithmic (dense, unpadded channels) FLOPs of the tcgen05 conv launches of one batch, as executed:
    the Head runs once per distinct frame (n_images); with the finest block's conv0[0] fused into the gather
    kernel its FLOPs are accounted to that kernel (returned separately)."""
    from vfi_b200.ifnet_layers import layer_info
    from vfi_b200.ifnet_params import BLOCK_C
    fl = 0
    for layer in (1, 2):  # Head cnn1/cnn2 at half resolution
        li = layer_info(layer)
        fl += 2 * n_images * (Hp // 2) * (Wp // 2) * 9 * li["cin"] * li["cout"]
    fl += 2 * n_images * Hp * Wp * 4 * 16 * 4  # Head cnn3: ConvTranspose 4x4 s2, 4 taps per output pixel
    fused = 0
    for b in range(5):
        s = 16 >> b
        hs, ws = Hp // s, Wp // s
        base = 4 + 11 * b
        n = B * T
        li = layer_info(base)
        f0 = 2 * n * (hs // 2) * (ws // 2) * 9 * li["cin"] * li["cout"]
        if b == 4 and fused_conv0a:
            fused = f0
        else:
            fl += f0
        li = layer_info(base + 1)
        fl += 2 * n * (hs // 4) * (ws // 4) * 9 * li["cin"] * li["cout"]
        c = BLOCK_C[b]
        fl += 8 * 2 * n * (hs // 4) * (ws // 4) * 9 * c * c
        # ConvTranspose 4x4 s2: 4 taps per output pixel; the final block computes (and is credited with) only the
        # 20 of 52 output channels the blend uses: 4 phases x (flow 4 + mask 1)
        fl += 2 * n * (hs // 2) * (ws // 2) * 4 * c * (20 if b == 4 else 52)
    return fl, fused


class ClockSampler(threading.Thread):
    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.idx = gpu_index
        self.rows = []
        self.stop = threading.Event()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop.is_set():
            try:
                o = subprocess.run(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) >= 6 and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def measure_aux(eng, clip_dev, H, W, hbm):
    """The path's other kernels (metrics + routing statistics): CUDA-event time per
    1080p frame on inputs larger than L2, against the HBM roofline with their algorithmic bytes."""
    from vfi_b200.engine import HudState
    n = min(24, clip_dev.shape[0] - 1)
    a, b = clip_dev[:n], clip_dev[1:n + 1]
    px = H * W
    flow = torch.randn((8, H, W, 2), device=clip_dev.device, dtype=torch.float32) * 4.0
    hud = HudState(eng)
    frames = [clip_dev[i] for i in range(n)]
    state = {"i": 0}

    def hud_step():
        hud.update(frames[state["i"] % n])
        state["i"] += 1
    cases = {
        "psnr_sse (A13)": (lambda: eng.psnr_sse(a, b), n, 6 * px, "both u8 RGB frames"),
        "ssim_full (A14)": (lambda: eng.ssim(a, b, 1), n, 6 * px, "both u8 RGB frames; gray + 7x7 windows on chip"),
        "ssim_quarter (A6 scene cut)": (lambda: eng.ssim(a, b, 4), n, 6 * px, "both u8 RGB frames"),
        "flow_stats (A7)": (lambda: eng.flow_stats(flow), 8, 8 * px, "fp32 flow field (u, v)"),
        "laplacian_var (A8)": (lambda: eng.laplacian_var(a), n, 3 * px, "one u8 RGB frame"),
        "hud_update (A9)": (hud_step, 1, 3 * px + 2 * px, "u8 RGB frame in, full-resolution mask out (+ 180x320 ring)"),
        "psnr_ssim_fused (A13+A14)": (lambda: eng.psnr_ssim(a, b), n, 6 * px, "both u8 RGB frames, read once for both metrics"),
        "farneback_flow_stats (N1+A7, 8 pairs/call)": (lambda: eng.farneback(a[:8], b[:8], want_flow=False, want_stats=True), 8, 6 * px,
                                                        "both u8 RGB frames (algorithmic input; the pyramid planes are the kernel's own traffic)"),
        "farneback_flow_stats (N1+A7, 1 pair/call)": (lambda: eng.farneback(a[:1], b[:1], want_flow=False, want_stats=True), 1, 6 * px,
                                                       "both u8 RGB frames"),
    }
    out = {}
    for name, (fn, per_call, bytes_per_frame, what) in cases.items():
        for _ in range(6):
            fn()
        torch.cuda.synchronize()
        reps = 10
        if per_call == 1 and name.startswith("hud"):
            # one small launch per call: back-to-back calls from Python are host-bound (~25 us of interpreter + allocator per
            # call), so the kernel itself is timed with the library's own events around its launch
            eng.profile_read()
            eng.profile(True)
            for _ in range(reps):
                fn()
            ms = eng.profile_read()["reduce"][0] / (reps * per_call)
            eng.profile(False)
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / (reps * per_call)
        gbps = bytes_per_frame / (ms / 1000.0) / 1e9
        out[name] = {"ms_per_frame": ms, "bound": "hbm", "achieved": gbps, "peak": hbm, "unit": "GB/s", "frac": gbps / hbm,
                     "algorithmic_bytes_per_frame": bytes_per_frame, "what": what}
    return out


def run_reference(args, rank, world):
    """The reference's own CPU path for this hot path = the oracle restatement (the reference's IFNet lives in a
    missing submodule; nothing of it can be installed).  One step = one Model.inference (1 pair, 1 timestep)."""
    if rank != 0:
        return
    from oracle import ifnet_ref
    import vfi_b200  # noqa: F401
    from vfi_b200 import ifnet_params
    from vfi_b200.synth import gen_clip
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    m = ifnet_ref.Model()
    m.flownet.load_state_dict(ifnet_params.random_init(1234), strict=True)
    fr = gen_clip(2, H, W, seed=1, cut_every=0, fast=None)
    i0 = ifnet_ref.pad_reflect(ifnet_ref.to_tensor(fr[0]), 64)
    i1 = ifnet_ref.pad_reflect(ifnet_ref.to_tensor(fr[1]), 64)

    def step(k):
        mid = m.inference(i0, i1, timestep=TIMESTEPS[k % 3], scale=1.0)[:, :, :H, :W]
        return ifnet_ref.to_numpy_u8(mid)

    for k in range(args.warmup):
        step(k)
    t0 = time.perf_counter()
    for k in range(args.steps):
        step(k)
    dt = time.perf_counter() - t0
    v = args.steps / dt
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1000, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "RIFE v4.25 4x, 1080p synthetic clip, CPU oracle (reference IFNet source is a missing "
                               "third-party submodule); one step = one inference (1 pair, 1 timestep)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} inferences of one 1080p pair, torch fp32, {cores} threads"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def pad_to(H, W, mult=64):
    return (H + mult - 1) // mult * mult, (W + mult - 1) // mult * mult


class Ctx:
    """Per-process state shared by the config runners."""
    def __init__(self, args):
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dev = torch.device("cuda", self.local)
        self.gloo = None

    def barrier(self):
        import torch.distributed as dist
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def cpu_wait(self):
        """Rendezvous on the HOST (gloo): ranks > 0 sleep here while rank 0 runs its single-rank extras, instead of spinning
        in an NCCL barrier kernel that would show up as GPU activity."""
        import torch.distributed as dist
        if self.world > 1:
            torch.cuda.synchronize()
            dist.barrier(group=self.gloo)

    def max_ms(self, ms):
        import torch.distributed as dist
        t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def timed(self, fn, steps, warmup, after=None):
        """W warm-up steps, barrier + synchronize, EXACTLY `steps` steps between two CUDA events on the launching stream,
        barrier + synchronize, max over ranks."""
        for k in range(warmup):
            fn(k)
        if after:
            after()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(steps):
            fn(warmup + k)
        if after:
            after()
        e1.record()
        self.barrier()
        return self.max_ms(e0.elapsed_time(e1))


def rife_roofline(eng, step, psteps, B, T, Hp, Wp, n_images, with_traffic):
    """Per-kernel-class device time over a profiled replay of `psteps` steps (CUDA events around every launch on the
    launching stream) -> (roofline dict for the tcgen05 conv class, breakdown dict)."""
    eng.profile(True)
    for k in range(psteps):
        step(k)
    prof = eng.profile_read()
    eng.profile(False)
    hbm, tf_burst, tf_sust, src = peaks()
    tot = sum(v[0] for v in prof.values())
    breakdown = {k: {"ms_per_step": v[0] / psteps, "launches_per_step": v[1] / psteps, "share": v[0] / tot if tot else 0}
                 for k, v in prof.items() if v[1]}
    tc_ms, tc_n = prof["conv_tcgen05"]
    if not tc_n:
        return None, breakdown
    fused_on = prof["gather_conv0a"][1] > 0
    fl, fl_fused = tc_flops_per_batch(B, T, Hp, Wp, n_images, fused_on)
    ach = fl * psteps / (tc_ms / 1000.0) / 1e12
    px = B * T * Hp * Wp   # padded full-resolution pixels of one step
    roof = {"kernel": "tcgen05 implicit-GEMM conv kernels (conv_v2 / conv_s2 / conv_tc: every conv launch of the step "
                      "except the one fused into the gather)",
            "bound": "tensor", "achieved": ach, "peak": tf_sust, "unit": "TFLOP/s", "frac": ach / tf_sust,
            "traffic": ncu_traffic(r"conv_(v2|s2|tc)_kernel") if with_traffic else None,
            "peak_source": f"{src} bf16_tflops_sustained", "launches": tc_n // psteps,
            "avg_launch_us": tc_ms / tc_n * 1000.0, "algorithmic_gflop_per_step": fl / 1e9,
            "flop_accounting": "executed work: Head once per distinct frame, final lastconv credited with the 20 of 52 "
                               "output channels it computes",
            "note": "N = 16..64 output channels per MMA: the A operand read from shared memory (4 KB per "
                    "128x16 MMA at 128 B/clk) bounds these layers at 25-50 % of the tensor pipe; see DESIGN.md"}
    others = {}

    def hbm_entry(key, bytes_per_px, what, pattern, extra=None):
        ms_k, n_k = prof[key]
        if n_k:
            by = px * bytes_per_px
            g = by * psteps / (ms_k / 1000.0) / 1e9
            others[key] = {"bound": "hbm", "achieved": g, "peak": hbm, "unit": "GB/s", "frac": g / hbm,
                           "algorithmic_bytes_per_px": bytes_per_px, "ms_per_step": ms_k / psteps, "what": what,
                           "traffic": ncu_traffic(pattern) if with_traffic else None}
            if extra:
                others[key].update(extra(ms_k / psteps))
    # algorithmic bytes per padded full-resolution pixel of one (pair, timestep) item
    hbm_entry("warp_blend", 19, "SURVEY 8(d): both u8 RGB frames 6, flow 4 + mask 1 as 16-bit 10, u8 out 3", r"final_blend_kernel",
              lambda ms: {"design_bytes_per_px": 43, "design_frac": px * 43 / (ms / 1000.0) / 1e9 / hbm,
                          "design_what": "what this design moves: RGBX taps of both frames 8, blocks 0-3 flow (fp32) 16, last "
                                         "block's flow delta + mask (fp16 texel) 16, u8 out 3"})
    hbm_entry("gather_conv0a", 8 + 16 + 8.5 + 1.5 + 16 + 8, "RGBX taps 8, Head feature taps 16, block-3 planes 8.5 "
              "(34 B / 4 px), coarser flows 1.5, summed flow out 16, conv0[0] output 8 (32 B / 4 px)", r"gather_conv0a(_p)?_kernel")
    hbm_entry("gather", (8 + 16 + 3 + 16) + (8 + 16 + 2 + 16) / 4.0 + (8 + 16 + 2 + 16) / 16.0 + 1,
              "blocks 0-3: taps + coarse inputs + 64 B per output pixel of the block input (s = 2: every pixel "
              "sampled; s = 4, 8, 16: 1/4, 1/16, 1/64 of them)", r"gather_(s2|lo|tile)?_?kernel")
    roof["other_kernels"] = others
    if fused_on:
        roof["fused_gather_conv0a_gflop_per_step"] = fl_fused / 1e9
    return roof, breakdown


# ---- BASELINE.json config 4: RIFE 4x at 1440p and 4K, 8 pairs per GPU per step ------------------------------------
def run_config4(cx, eng, prec):
    from vfi_b200.synth import gen_clip
    out = {}
    B, T = 8, len(TIMESTEPS)
    for name, (h, w), seed in (("1440p", (1440, 2560), 3), ("2160p", (2160, 3840), 4)):
        clip = torch.from_numpy(gen_clip(B, h, w, seed=seed * 100 + cx.rank, cut_every=0, fast=None)).to(cx.dev)
        dst = torch.empty((B, T, h, w, 3), dtype=torch.uint8, device=cx.dev)

        def step(k):
            eng.infer(clip[:-1], clip[1:], TIMESTEPS, 1.0, prec, out=dst)
        steps = 6
        ms = cx.timed(step, steps, 3)
        fps = cx.world * B * T * steps / (ms / 1000.0)
        entry = {"frames_per_s": fps, "ms_per_step": ms / steps, "pairs_per_step_per_gpu": B, "resolution": [h, w],
                 "padded": list(pad_to(h, w)), "steps": steps, "warmup": 3,
                 "l2_policy": "one step's activations (several GB) exceed the 126 MB L2"}
        if cx.rank == 0:
            Hp, Wp = pad_to(h, w)
            roof, br = rife_roofline(eng, step, 3, B, T, Hp, Wp, B + 1, False)
            if roof:
                entry["roofline"] = {"bound": "tensor", "achieved": roof["achieved"], "peak": roof["peak"], "unit": "TFLOP/s",
                                     "frac": roof["frac"], "algorithmic_gflop_per_step": roof["algorithmic_gflop_per_step"],
                                     "gflop_per_frame_survey_8d": {"1440p": 152.22, "2160p": 337.52}[name]}
                wb = roof["other_kernels"].get("warp_blend")
                if wb:
                    entry["warp_blend"] = {k: wb[k] for k in ("achieved", "frac", "unit", "algorithmic_bytes_per_px", "design_frac", "ms_per_step")}
            entry["kernel_ms_per_step"] = {k: round(v["ms_per_step"], 4) for k, v in br.items()}
        out[name] = entry
        del clip, dst
        torch.cuda.empty_cache()
    return out


# ---- BASELINE.json config 3: AdaptiveVFI routing + RIFE fast path over the 300-pair clip, sharded by pair index ----
CONFIG3 = dict(n_pairs=300, H=1080, W=1920, seed=2, cut_every=60, fast=(120, 150))


def run_config3(cx, group):
    import contextlib
    from vfi_b200 import sharding
    from vfi_b200.models.novel.adaptive_pipeline import AdaptivePipeline
    from vfi_b200.synth import gen_clip
    n = CONFIG3["n_pairs"]
    lo, hi = sharding.shard_range(n, cx.rank, cx.world)
    base = max(0, lo - 16)
    t0 = time.perf_counter()
    host = gen_clip(hi - base, CONFIG3["H"], CONFIG3["W"], seed=CONFIG3["seed"], cut_every=CONFIG3["cut_every"],
                    fast=CONFIG3["fast"], start=base)
    gen_s = time.perf_counter() - t0
    frames = torch.from_numpy(host).to(cx.dev)
    pipe = AdaptivePipeline(device=f"cuda:{cx.local}", enable_vfimamba=False, allow_random_weights=True)
    with contextlib.redirect_stdout(sys.stderr):
        pipe.load()
    dst = torch.empty((hi - lo, 3, CONFIG3["H"], CONFIG3["W"], 3), dtype=torch.uint8, device=cx.dev)
    res = {}

    def run(k, collect=True):
        pipe.router.reset()
        pipe.reset_stats()
        res["r"] = pipe.process_clip_batched(frames, 3, cx.rank, cx.world, group, batch=cx.args.adaptive_batch, out=dst,
                                             n_pairs=n, base=base, collect=collect)
    eng = pipe.router.engine
    l0 = eng.launches
    ms = cx.timed(run, 1, 1)
    launches = (eng.launches - l0) // 2
    r = res["r"]
    entry = {"pairs_per_s": n / (ms / 1000.0), "ms_per_clip": ms, "pairs": n, "frames_out": n * 3,
             "workload": "AdaptivePipeline.process_clip_batched over the SURVEY 8(d) config-3 clip (1080p, a cut every 60 pairs, "
                         "fast segment 120-150, static HUD), %d pairs per batch, clip resident in HBM, results resident in HBM; "
                         "per pair: scene-cut SSIM, Farneback + flow statistics, Laplacian variance, HUD update, decision tree, "
                         "RIFE 4x with the HUD composite; one all-reduce of the route counters + all-gather of routes / coverage "
                         "inside the timed region" % cx.args.adaptive_batch,
             "shard": [lo, hi], "clip_gen_s_rank0": round(gen_s, 2), "gpu_launches_per_clip_rank0": int(launches),
             "stats": r["stats"]}
    if cx.rank == 0:
        gpath = os.path.join(ROOT, "tests", "golden", "config3_1080p.npz")
        if os.path.exists(gpath):
            g = np.load(gpath)
            entry["parity"] = {"golden": "tests/golden/config3_1080p.npz (the reference's AdaptiveRouter.analyze run serially, cv2 Farneback)",
                               "routes_bit_exact": bool(np.array_equal(r["routes"], g["routes"])),
                               "cuts_bit_exact": bool(np.array_equal(r["cuts"].astype(np.int8), g["cuts"])),
                               "hud_coverage_equal": bool(np.array_equal(r["hud_coverage"], g["hud_coverage"])),
                               "n_ranks": cx.world}
        # where the clip's time goes on one rank: the analysis kernels against the interpolation
        r_keep = res["r"]
        eng.profile(True)
        run(0, collect=False)            # rank 0 alone: no collective in this replay
        prof = eng.profile_read()
        eng.profile(False)
        res["r"] = r_keep
        entry["kernel_ms_per_shard_rank0"] = {k: round(v[0], 3) for k, v in prof.items() if v[1]}
    del frames, dst, pipe
    torch.cuda.empty_cache()
    return entry, host


# ---- BASELINE.json config 5: PSNR / SSIM of 10 000 1080p frame pairs, sharded by frame index ----------------------
def run_config5(cx, group, n_total):
    from vfi_b200 import sharding
    from vfi_b200.engine import Engine
    from vfi_b200.evaluation import metrics as M
    from vfi_b200.synth import gen_clip
    eng = Engine.get(cx.local)
    CH = 50
    lo, hi = sharding.shard_range(n_total, cx.rank, cx.world)
    gt_clip = torch.from_numpy(gen_clip(CH - 1, H, W, seed=5, cut_every=0, fast=None)).to(cx.dev)   # 50 ground-truth frames
    n_gt = CH
    pred = torch.empty((hi - lo, H, W, 3), dtype=torch.uint8, device=cx.dev)
    gen = torch.Generator(device=cx.dev)
    exact_sse = 0
    for a in range(lo, hi, CH):            # pred = clip(gt + U{-3..3}), generated on the device, seeded per 50-frame chunk
        b = min(hi, a + CH)
        gen.manual_seed(5 * 1000003 + a // CH)
        g0 = (a % n_gt)
        gt = gt_clip[g0:g0 + (b - a)].to(torch.int16)
        noise = torch.randint(-3, 4, gt.shape, generator=gen, device=cx.dev, dtype=torch.int16)
        p = (gt + noise).clamp_(0, 255)
        exact_sse += int(((p - gt).to(torch.int64) ** 2).sum())
        pred[a - lo:b - lo] = p.to(torch.uint8)
    del gt, noise, p
    torch.cuda.synchronize()
    gt_fn = lambda a, b: gt_clip[a % n_gt:a % n_gt + (b - a)]
    pred_fn = lambda a, b: pred[a - lo:b - lo]
    out = {}

    def run(k):
        part = M.evaluate_range(pred_fn, gt_fn, lo, hi, CH, eng)
        out["part"] = part
        out["res"] = M.reduce_sums(part, (lo, hi), group, cx.dev)
    l0 = eng.launches
    ms = cx.timed(run, 1, 1)
    launches = (eng.launches - l0) // 2
    res = out["res"]
    import torch.distributed as dist
    ex = torch.tensor([exact_sse], dtype=torch.int64, device=cx.dev)
    parts = torch.tensor(out["part"], dtype=torch.float64, device=cx.dev)
    if cx.world > 1:                      # verification collectives (outside the timed region)
        dist.all_reduce(ex, group=group)
        gathered = [torch.empty_like(parts) for _ in range(cx.world)]
        dist.all_gather(gathered, parts, group=group)
    else:
        gathered = [parts]
    serial = torch.stack(gathered).cpu().numpy()
    s_psnr, s_ssim = float(np.sum(serial[:, 0])), float(np.sum(serial[:, 1]))
    px = H * W
    hbm = peaks()[0]
    gbps = n_total * 6 * px / (ms / 1000.0) / 1e9
    entry = {"frame_pairs_per_s": n_total / (ms / 1000.0), "ms_total": ms, "frame_pairs": n_total, "psnr_mean": res["psnr"],
             "ssim_mean": res["ssim"], "gpu_launches_rank0": int(launches),
             "workload": "vfi_psnr_ssim_u8 (both metrics from one read) over 10 000 1080p frame pairs resident in HBM, pred = "
                         "clip(gt + U{-3..3}) generated on the device (seeded per 50-frame chunk), gt = a 50-frame synthetic clip "
                         "(seed 5; 311 MB per chunk > L2); one all-reduce of {sum psnr, sum ssim, n, n_inf, sum sse} inside the timed region",
             "roofline": {"bound": "hbm", "achieved": gbps / cx.world, "peak": hbm, "unit": "GB/s per GPU", "frac": gbps / cx.world / hbm,
                          "algorithmic_bytes_per_pair": 6 * px},
             "parity": {"sse_sum_exact": bool(res["sse_sum"] == int(ex[0])), "n": res["n"] == n_total,
                        "psnr_sum_rel_err_vs_rank_order_sum": abs(res["psnr_sum"] - s_psnr) / max(1.0, abs(s_psnr)),
                        "ssim_sum_rel_err_vs_rank_order_sum": abs(res["ssim_sum"] - s_ssim) / max(1.0, abs(s_ssim)),
                        "n_ranks": cx.world}}
    del pred, gt_clip
    torch.cuda.empty_cache()
    return entry


# ---- BASELINE.json config 1 + the reference's call shape (one pair per call) --------------------------------------
def run_config1_and_latency(model_bf16, clip):
    """Rank 0 only.  Config 1: one 1080p pair, 2x (t = 0.5), fp32 mode.  Latency: RIFEModel.interpolate (the call
    scripts/run_benchmarks.py:142-149 times), host numpy frames in and out, one pair per call."""
    import contextlib
    from vfi_b200.models.sota.rife_wrapper import RIFEModel
    f0, f1 = clip[0], clip[1]
    out = {}
    with contextlib.redirect_stdout(sys.stderr):
        m32 = RIFEModel(device=str(model_bf16.device), precision="fp32", allow_random_weights=True)
        m32.load()
    d0, d1 = torch.from_numpy(f0).to(model_bf16.device)[None], torch.from_numpy(f1).to(model_bf16.device)[None]
    from vfi_b200 import _abi
    for name, m, prec in (("fp32", m32, _abi.PRECISION_FP32), ("bf16", model_bf16, _abi.PRECISION_BF16)):
        eng = m._engine
        for _ in range(3):
            eng.infer(d0, d1, [0.5], 1.0, prec)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5 if name == "fp32" else 20
        e0.record()
        for _ in range(reps):
            eng.infer(d0, d1, [0.5], 1.0, prec)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        out[f"config1_{name}"] = {"ms_per_pair": ms, "frames_per_s": 1000.0 / ms, "gflop_per_inference_survey_8d": 92.83,
                                  "achieved_tflops": 92.83 / ms, "workload": f"one 1080p pair, t = 0.5 (2x), {name} mode, frames resident"}
    lat = {}
    for T in (1, 3):
        ts = []
        for k in range(40):
            t0 = time.perf_counter()
            model_bf16.interpolate(f0, f1, T)
            ts.append((time.perf_counter() - t0) * 1000.0)
        ts = np.array(ts[8:])
        lat[f"num_frames_{T}"] = {"p50_ms": float(np.percentile(ts, 50)), "p99_ms": float(np.percentile(ts, 99)), "mean_ms": float(ts.mean())}
    out["latency_1pair"] = {"api": "RIFEModel.interpolate(frame0, frame1, num_frames) with host numpy uint8 frames in and out "
                                   "(pinned staging + H2D + kernels + D2H + synchronise), bf16, one pair per call",
                            "reference_budget_ms_per_pair": 33.33, **lat}
    return out


def cpu_baselines(clip, cores):
    """Rank 0 only, bounded samples on the host cores: IFNet oracle (the headline's baseline), the reference router's
    analyze() semantics (oracle/router_ref.py: cv2 + restated SSIM), and compute_psnr_simple + restated compute_ssim_simple."""
    import cv2
    from oracle import ifnet_ref, metrics_ref, router_ref
    from vfi_b200 import ifnet_params
    from vfi_b200.synth import noisy_copy
    torch.set_num_threads(cores)
    cv2.setNumThreads(cores)
    m = ifnet_ref.Model()
    m.flownet.load_state_dict(ifnet_params.random_init(1234), strict=True)
    i0 = ifnet_ref.pad_reflect(ifnet_ref.to_tensor(clip[0]), 64)
    i1 = ifnet_ref.pad_reflect(ifnet_ref.to_tensor(clip[1]), 64)
    m.inference(i0, i1, 0.5, 1.0)
    n = 0
    t0 = time.perf_counter()
    while n < 6 or (time.perf_counter() - t0 < 10 and n < 24):
        ifnet_ref.to_numpy_u8(m.inference(i0, i1, TIMESTEPS[n % 3], 1.0)[:, :, :H, :W])
        n += 1
    dt = time.perf_counter() - t0
    base = {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} oracle inferences (1080p, fp32, {cores} torch threads) of the same clip"}
    r = router_ref.RouterRef()
    k = 0
    t0 = time.perf_counter()
    while k < 4 or (time.perf_counter() - t0 < 8 and k < min(24, len(clip) - 1)):
        r.analyze(clip[k], clip[k + 1])
        k += 1
    dt = time.perf_counter() - t0
    base["routing"] = {"value": k / dt, "unit": "pairs/s", "cores": cores, "kind": "port",
                       "sample": f"{k} pairs of the config-3 clip through the reference's analyze() semantics (cv2 gray / resize / "
                                 f"Farneback / Laplacian / morphology with cv2.setNumThreads({cores}), SSIM restated on scipy)"}
    pred = noisy_copy(clip[:8], seed=5)
    k = 0
    t0 = time.perf_counter()
    while k < 4 or (time.perf_counter() - t0 < 6 and k < 200):
        metrics_ref.psnr_simple(pred[k % 8], clip[k % 8])
        metrics_ref.ssim_simple(pred[k % 8], clip[k % 8])
        k += 1
    dt = time.perf_counter() - t0
    base["metrics"] = {"value": k / dt, "unit": "frame pairs/s", "cores": cores, "kind": "port",
                       "sample": f"{k} 1080p frame pairs through compute_psnr_simple (verbatim) + compute_ssim_simple (skimage restated on scipy)"}
    return base


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--pairs", type=int, default=30, help="frame pairs per step (batch): the 300-pair clip in 10 steps; "
                    "the per-launch fixed cost of the ~67 launches of a step is amortised over more tiles (8 -> 30 pairs: +7 %)")
    ap.add_argument("--pool", type=int, default=60, help="unique synthetic pairs resident in HBM")
    ap.add_argument("--adaptive-batch", type=int, default=int(os.environ.get("VFI_ADAPTIVE_BATCH", "16")), help="config 3: pairs per batch")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-aux", action="store_true", help="skip timing the metric / routing-statistics kernels")
    ap.add_argument("--no-other-configs", action="store_true", help="headline (config 2) only")
    ap.add_argument("--eval-frames", type=int, default=10000, help="config 5: frame pairs over all ranks")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    args = ap.parse_args()

    cx = Ctx(args)
    rank, world, local = cx.rank, cx.world, cx.local
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch.distributed as dist
    import vfi_b200  # noqa: F401
    from vfi_b200 import _abi
    from vfi_b200.models.sota.rife_wrapper import RIFEModel
    from vfi_b200.synth import gen_clip

    torch.cuda.set_device(local)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version banner (and, with NCCL_DEBUG set, its log) to file descriptor 1 when the communicator is
        # created: point fd 1 at stderr for that moment so that stdout carries exactly the one JSON line
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            import datetime
            dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=180))
            dist.barrier()      # forces communicator creation (eager with device_id, but be sure)
            torch.cuda.synchronize()
            cx.gloo = dist.new_group(backend="gloo")
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    dev = cx.dev

    model = RIFEModel(device=f"cuda:{local}", precision=args.precision, allow_random_weights=True)
    import contextlib
    with contextlib.redirect_stdout(sys.stderr):   # the wrapper prints "Loading RIFE ..." like the reference; stdout is the JSON line only
        model.load()
    eng = model._engine
    for opt in ("l2_chunk_mb", "fuse_gather", "conv_v2", "v2_tma", "v2_res_global", "v2_pm", "gc_ablate", "gc_persistent", "fuse_gather3", "t_fast"):  # tuning knobs of the library (defaults are the shipped ones)
        if os.environ.get("VFI_" + opt.upper()) is not None:
            eng.set_option(opt, int(os.environ["VFI_" + opt.upper()]))
    B, T = args.pairs, len(TIMESTEPS)
    # each rank works on its own shard of the clip (frame-pair sharding, no data-path collective)
    clip = gen_clip(args.pool, H, W, seed=1 + rank, cut_every=0, fast=None)
    clip_dev = torch.from_numpy(clip).to(dev)      # the clip resident in HBM: pair k = frames k, k+1
    out_dev = torch.empty((B, T, H, W, 3), dtype=torch.uint8, device=dev)
    prec = _abi.PRECISION_BF16 if args.precision == "bf16" else _abi.PRECISION_FP32

    def batch_idx(k):
        s = (k * B) % (args.pool - B + 1) if args.pool > B else 0
        return slice(s, s + B)

    def step_resident(k):
        sl = batch_idx(k)
        eng.infer(clip_dev[sl.start:sl.stop], clip_dev[sl.start + 1:sl.stop + 1], TIMESTEPS, 1.0, prec, out=out_dev)

    clip_host = torch.from_numpy(clip).pin_memory()
    e2e_depth = int(os.environ.get("VFI_E2E_DEPTH", "2"))   # buffer slots of the pipelined public API
    streamer = model.clip_streamer(B, H, W, TIMESTEPS, depth=e2e_depth)

    def step_e2e(k):
        # public pipelined API: B+1 consecutive pinned host frames in (uint8), B x T pinned host frames out, every step
        sl = batch_idx(k)
        streamer.submit(clip_host[sl.start:sl.stop + 1])

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    warm = max(args.warmup, 3)
    l0 = eng.launches
    ms = cx.timed(step_resident, args.steps, warm)
    launches = (eng.launches - l0) * args.steps // (args.steps + warm)
    if rank == 0:
        sampler.stop.set()
    frames_per_step = B * T
    value = world * frames_per_step * args.steps / (ms / 1000.0)

    e2e_steps = max(3, min(args.steps, 25))
    ms_e2e = cx.timed(step_e2e, e2e_steps, 3, after=streamer.join)   # ends when the last result has reached pinned host memory
    e2e = world * frames_per_step * e2e_steps / (ms_e2e / 1000.0)
    del streamer
    # the same leg with the optional planar YUV 4:2:0 result format (half the device -> host bytes): secondary figure
    streamer = model.clip_streamer(B, H, W, TIMESTEPS, depth=e2e_depth, output="yuv420")
    ms_yuv = cx.timed(step_e2e, e2e_steps, 3, after=streamer.join)
    e2e_yuv = world * frames_per_step * e2e_steps / (ms_yuv / 1000.0)
    del streamer

    # ---- the other BASELINE.json configs, sharded over the same ranks ----
    other = {}
    clip3 = None
    if not args.no_other_configs and args.precision == "bf16":
        other["config4_rife4x_1440p_2160p"] = run_config4(cx, eng, prec)
        other["config3_adaptive_1080p_300pairs"], clip3 = run_config3(cx, group)
        other["config5_psnr_ssim_10k_1080p"] = run_config5(cx, group, args.eval_frames)

    # ---- rank 0 alone from here; the other ranks sleep on a host-side barrier ----
    roof = breakdown = aux = cpu_base = None
    if rank == 0:
        if args.precision == "bf16":
            roof, breakdown = rife_roofline(eng, step_resident, min(args.steps, 10), B, T, 1088, 1920, B + 1, True)
        else:
            _, breakdown = rife_roofline(eng, step_resident, min(args.steps, 3), B, T, 1088, 1920, B + 1, False)
        if not args.no_aux:
            aux = measure_aux(eng, clip_dev, H, W, peaks()[0])
        if not args.no_other_configs and args.precision == "bf16":
            other.update(run_config1_and_latency(model, clip))
        if not args.no_cpu_baseline:
            cpu_base = cpu_baselines(clip3 if clip3 is not None else clip, os.cpu_count() or 1)
    cx.cpu_wait()

    if rank == 0:
        sampler.join(timeout=2)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": {"workload": "RIFE v4.25 4x (t=0.25/0.5/0.75) on a synthetic 1080p clip, random-init weights "
                                   "(BASELINE.json configs[1])",
                       "pairs_per_step": B, "frames_per_step": frames_per_step, "pool_pairs": args.pool,
                       "pool_note": "the clip is walked as batches of consecutive pairs rotating through a pool of distinct resident pairs "
                                    "(same convs; the gather kernels' cache behaviour follows the pool's flow fields)",
                       "resolution": [H, W], "padded": [1088, 1920], "parallelism": f"frame-pair sharding x{world}",
                       "l2_policy": "inputs+activations per step (>1 GB) exceed the 126 MB L2; batches rotate through the pool"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int((B + 1) * H * W * 3),
                    "d2h_bytes_per_step": int(B * T * H * W * 3), "steps": e2e_steps,
                    "yuv420_output": {"value": e2e_yuv, "unit": UNIT, "d2h_bytes_per_step": int(B * T * H * W * 3 // 2),
                                      "what": "same call with clip_streamer(output='yuv420'): results come down as cv2-exact I420 frames "
                                              "(1.5 B/px); the RGB figure above is the headline (the reference's result format)"},
                    "api": "RIFEModel.clip_streamer().submit (B+1 pinned host uint8 frames in, B x T pinned host uint8 frames out per step; upload / kernels / download on three streams, %d slots)" % e2e_depth},
            "gpu_launches": int(launches),
            "clocks": sampler.summary(),
            "roofline": roof, "kernel_breakdown": breakdown, "aux_kernels": aux, "cpu_baseline": cpu_base,
            "other_configs": other or None,
        }
        print(json.dumps(line))
    if world > 1:
        cx.cpu_wait()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
