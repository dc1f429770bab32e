#!/usr/bin/env python
"""bench.py -- interpolated 1080p frames/s, RIFE v4.25 4x (t = 0.25/0.5/0.75), bf16, synthetic clip.

    python bench.py [--gpus N] [--steps K] [--warmup W]            (N > 1: launched under torchrun)
    python bench.py --impl reference ...                           (CPU oracle on the host cores)

One "step" = one pass of the hot path over one batch of `--pairs` frame pairs x 3 timesteps
(BASELINE.json configs[1]; K*pairs = 300 pairs by default).  `value` is device-timed with the clip
resident in HBM; `e2e` goes through RIFEModel.interpolate_batch with pinned HOST buffers (H2D + D2H
inside the timed region).  Prints ONE JSON line on rank 0.
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


def ncu_conv_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per tcgen05 conv launch (bytes, average over the step's conv launches) from
    the committed ncu launch list of this workload, profiles/r01_launches_final.txt; None if that file is absent."""
    import re
    path = os.path.join(ROOT, "profiles", "r01_launches_final.txt")
    if not os.path.exists(path):
        return None
    tot, n = 0.0, 0
    for ln in open(path):
        if ln.startswith("#") or not re.search(r"conv_(v2|s2|tc)_kernel", ln):
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
        # ConvTranspose 4x4 s2: 4 taps per output pixel; the final block computes only the 20 of 52 channels used
        fl += 2 * n * (hs // 2) * (ws // 2) * 4 * c * 52
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=75)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--pairs", type=int, default=8, help="frame pairs per step (batch)")
    ap.add_argument("--pool", type=int, default=24, help="unique synthetic pairs resident in HBM")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-aux", action="store_true", help="skip timing the metric / routing-statistics kernels")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch.distributed as dist
    import vfi_b200  # noqa: F401
    from vfi_b200 import _abi
    from vfi_b200.models.sota.rife_wrapper import RIFEModel
    from vfi_b200.synth import gen_clip

    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version banner (and, with NCCL_DEBUG set, its log) to file descriptor 1 when the communicator is
        # created: point fd 1 at stderr for that moment so that stdout carries exactly the one JSON line
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()      # forces communicator creation (eager with device_id, but be sure)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    dev = torch.device("cuda", local)

    model = RIFEModel(device=f"cuda:{local}", precision=args.precision, allow_random_weights=True)
    import contextlib
    with contextlib.redirect_stdout(sys.stderr):   # the wrapper prints "Loading RIFE ..." like the reference; stdout is the JSON line only
        model.load()
    eng = model._engine
    for opt in ("l2_chunk_mb", "fuse_gather", "conv_v2", "v2_tma", "v2_res_global", "v2_pm"):  # tuning knobs of the library (defaults are the shipped ones)
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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for k in range(warmup):
            fn(k)
        if fn is step_e2e:
            streamer.join()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = eng.launches
        e0.record()
        for k in range(steps):
            fn(warmup + k)
        if fn is step_e2e:
            streamer.join()     # the timed region ends when the last result has reached pinned host memory
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0]), eng.launches - l0

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms, launches = timed(step_resident, args.steps, max(args.warmup, 3))
    if rank == 0:
        sampler.stop.set()
    frames_per_step = B * T
    value = world * frames_per_step * args.steps / (ms / 1000.0)

    e2e_steps = max(3, min(args.steps, 25))
    ms_e2e, _ = timed(step_e2e, e2e_steps, 3)
    e2e = world * frames_per_step * e2e_steps / (ms_e2e / 1000.0)

    # per-kernel-class device time over a profiled replay of the same steps (events around every launch)
    roof = None
    breakdown = None
    if rank == 0:
        eng.profile(True)
        psteps = min(args.steps, 10)
        for k in range(psteps):
            step_resident(k)
        prof = eng.profile_read()
        eng.profile(False)
        hbm, tf_burst, tf_sust, src = peaks()
        tot = sum(v[0] for v in prof.values())
        breakdown = {k: {"ms_per_step": v[0] / psteps, "launches_per_step": v[1] / psteps,
                         "share": v[0] / tot if tot else 0} for k, v in prof.items() if v[1]}
        tc_ms, tc_n = prof["conv_tcgen05"]
        if tc_n and args.precision == "bf16":
            fused_on = prof["gather_conv0a"][1] > 0
            fl, fl_fused = tc_flops_per_batch(B, T, 1088, 1920, B + 1, fused_on)
            ach = fl * psteps / (tc_ms / 1000.0) / 1e12
            px = B * T * 1088 * 1920   # padded full-resolution pixels of one step
            roof = {"kernel": "tcgen05 implicit-GEMM conv kernels (conv_v2 / conv_s2 / conv_tc: every conv launch of the step "
                              "except the one fused into the gather)",
                    "bound": "tensor", "achieved": ach, "peak": tf_sust, "unit": "TFLOP/s", "frac": ach / tf_sust,
                    "traffic": ncu_conv_traffic(), "peak_source": f"{src} bf16_tflops_sustained", "launches": tc_n // psteps,
                    "avg_launch_us": tc_ms / tc_n * 1000.0, "algorithmic_gflop_per_step": fl / 1e9,
                    "note": "N = 16..64 output channels per MMA: the A operand read from shared memory (4 KB per "
                            "128x16 MMA at 128 B/clk) bounds these layers at 25-50 % of the tensor pipe; see DESIGN.md"}
            others = {}

            def hbm_entry(key, bytes_per_px, what):
                ms_k, n_k = prof[key]
                if n_k:
                    by = px * bytes_per_px
                    others[key] = {"bound": "hbm", "achieved": by * psteps / (ms_k / 1000.0) / 1e9, "peak": hbm, "unit": "GB/s",
                                   "frac": by * psteps / (ms_k / 1000.0) / 1e9 / hbm, "algorithmic_bytes_per_px": bytes_per_px,
                                   "ms_per_step": ms_k / psteps, "what": what}
            # algorithmic bytes per padded full-resolution pixel of one (pair, timestep) item
            hbm_entry("warp_blend", 8 + 16 + 16 + 3, "RGBX taps of both frames 8, blocks 0-3 flow (fp32) 16, last block's "
                                                     "flow delta + mask (fp16 texel) 16, u8 out 3")
            hbm_entry("gather_conv0a", 8 + 16 + 8.5 + 1.5 + 16 + 8, "RGBX taps 8, Head feature taps 16, block-3 planes 8.5 "
                      "(34 B / 4 px), coarser flows 1.5, summed flow out 16, conv0[0] output 8 (32 B / 4 px)")
            hbm_entry("gather", (8 + 16 + 3 + 16) + (8 + 16 + 2 + 16) / 4.0 + (8 + 16 + 2 + 16) / 16.0 + 1,
                      "blocks 0-3: taps + coarse inputs + 64 B per output pixel of the block input (s = 2: every pixel "
                      "sampled; s = 4, 8, 16: 1/4, 1/16, 1/64 of them)")
            roof["other_kernels"] = others
            if fused_on:
                roof["fused_gather_conv0a_gflop_per_step"] = fl_fused / 1e9

    aux = None
    if rank == 0 and not args.no_aux:
        aux = measure_aux(eng, clip_dev, H, W, peaks()[0])

    cpu_base = None
    if rank == 0 and not args.no_cpu_baseline:
        from oracle import ifnet_ref
        from vfi_b200 import ifnet_params
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
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
        cpu_base = {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
                    "sample": f"{n} oracle inferences (1080p, fp32, {cores} torch threads) of the same clip"}

    if rank == 0:
        sampler.join(timeout=2)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": {"workload": "RIFE v4.25 4x (t=0.25/0.5/0.75) on a synthetic 1080p clip, random-init weights "
                                   "(BASELINE.json configs[1])",
                       "pairs_per_step": B, "frames_per_step": frames_per_step, "pool_pairs": args.pool,
                       "resolution": [H, W], "padded": [1088, 1920], "parallelism": f"frame-pair sharding x{world}",
                       "l2_policy": "inputs+activations per step (>1 GB) exceed the 126 MB L2; batches rotate through the pool"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(streamer.h2d_bytes),
                    "d2h_bytes_per_step": int(B * T * H * W * 3), "steps": e2e_steps,
                    "api": "RIFEModel.clip_streamer().submit (B+1 pinned host uint8 frames in, B x T pinned host uint8 frames out per step; upload / kernels / download on three streams, %d slots)" % e2e_depth},
            "gpu_launches": int(launches),
            "clocks": sampler.summary(),
            "roofline": roof, "kernel_breakdown": breakdown, "aux_kernels": aux, "cpu_baseline": cpu_base,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
