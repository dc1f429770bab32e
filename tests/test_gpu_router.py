"""AdaptiveRouter / AdaptivePipeline / RIFEModel / metrics behind the reference interface, on the GPU:
routing decisions and scene-cut flags bit-exact against the reference-generated golden rows."""
import numpy as np
import pytest
import torch

import vfi_b200  # noqa: F401
from oracle import ifnet_ref, metrics_ref, router_ref
from vfi_b200.evaluation.metrics import QualityEvaluator, compute_psnr_simple, compute_ssim_simple, psnr_batch
from vfi_b200.models.novel.adaptive_pipeline import AdaptivePipeline, AdaptiveRouter, cv2_farneback_flow
from vfi_b200.models.sota.rife_wrapper import RIFEModel
from vfi_b200.scripts.run_benchmarks import get_available_models, run_model_benchmark
from vfi_b200.synth import gen_clip, noisy_copy

pytestmark = pytest.mark.gpu
CODE = {"rife": 0, "vfimamba": 1, "scene_change": 2}


@pytest.fixture(scope="module")
def clip_and_gold(golden):
    g = golden("router.npz")
    n, H, W, seed = g["clip"].tolist()
    return gen_clip(n, H, W, seed=seed, cut_every=int(g["cut_every"]), fast=tuple(g["fast"].tolist())), g


def _run(router, clip):
    rows, masks = [], []
    for i in range(len(clip) - 1):
        a = router.analyze(clip[i], clip[i + 1])
        rows.append([a.motion_mean, a.motion_max, a.motion_std, float(a.has_particles), float(a.is_scene_change),
                     a.hud_coverage, CODE[a.recommended_model], a.confidence])
        m = router.hud_mask_host()
        live = m is not None and not a.is_scene_change and len(router.frame_history) >= 5
        masks.append(np.packbits(m if live else np.zeros(clip[i].shape[:2], bool)))
    return np.array(rows), np.stack(masks)


@pytest.mark.parametrize("flow", ["device", "cv2"])
@pytest.mark.parametrize("tight", [False, True])
def test_router_vs_reference_golden(engine, clip_and_gold, tight, flow):
    """flow="device": the product path (CUDA Farneback, row N1) -- every decision bit-exact, the flow statistics within the
    flow tolerance.  flow="cv2": OpenCV's own flow uploaded, isolating the reduction kernels (max exact, mean/std 1e-5)."""
    clip, g = clip_and_gold
    kw = {}
    if tight:
        lo, hi, pt = g["tight"].tolist()
        kw = dict(motion_threshold_low=lo, motion_threshold_high=hi, particle_threshold=pt)
    rows, masks = _run(AdaptiveRouter(flow_fn=cv2_farneback_flow if flow == "cv2" else None, **kw), clip)
    gold = g["rows_tight" if tight else "rows"]
    for col in (3, 4, 6):                      # has_particles, is_scene_change, route: BIT-EXACT
        assert np.array_equal(rows[:, col], gold[:, col])
    assert np.array_equal(rows[:, 5], gold[:, 5])                       # HUD coverage: integer count / pixels
    if flow == "cv2":
        assert np.array_equal(rows[:, 1], gold[:, 1])                   # max |flow|: exact in any order
        assert np.allclose(rows[:, [0, 2]], gold[:, [0, 2]], rtol=1e-5, atol=1e-12)
        assert np.allclose(rows[:, 7], gold[:, 7], rtol=1e-5, atol=1e-9)
    else:
        assert np.allclose(rows[:, [0, 1, 2]], gold[:, [0, 1, 2]], rtol=2e-3, atol=2e-3)
        assert np.allclose(rows[:, 7], gold[:, 7], rtol=2e-3, atol=2e-3)
    if not tight:
        assert np.array_equal(masks, g["masks"])


def test_router_public_methods(engine, clip_and_gold):
    clip, g = clip_and_gold
    r, ref = AdaptiveRouter(), router_ref.RouterRef()
    cut, score = r.detect_scene_change(clip[5], clip[6])
    assert cut and abs(score - router_ref.scene_ssim(clip[5], clip[6])) < 1e-9
    mean, mx, sd, mag = r.compute_motion(clip[0], clip[1])             # CUDA Farneback: flow tolerance, not bit-exact
    assert abs(mx - g["motion"][0][1]) < 2e-3 and abs(mean - g["motion"][0][0]) < 1e-4 and mag.shape == clip[0].shape[:2]
    mean_c, mx_c, _, _ = AdaptiveRouter(flow_fn=cv2_farneback_flow).compute_motion(clip[0], clip[1])
    assert mx_c == g["motion"][0][1] and abs(mean_c - g["motion"][0][0]) < 1e-5
    hp, ps = r.detect_particles(clip[0], mag, sd)
    want = float(np.sqrt(min(sd / 20.0, 1.0) * min(g["laplacian_var"][0] / 500.0, 1.0)))
    assert hp == (want > 0.4) and abs(ps - want) < 1e-12
    for f in clip[:6]:
        m, c = r.detect_hud(f)
        m2, c2 = ref.detect_hud(f)
        assert np.array_equal(m, m2) and c == c2 and m.dtype == bool
    assert len(r.frame_history) == 6 and np.array_equal(r.frame_history[-1], ref.frame_history[-1])


def test_pipeline_routes_composites_and_counts(engine, clip_and_gold, oracle_model):
    clip, g = clip_and_gold
    p = get_available_models()["adaptive"]()
    p.load()
    assert p.enable_vfimamba is False                                   # same fallback as the reference (:358-361)
    ref = router_ref.RouterRef()
    for i in range(8):
        mids = p.interpolate(clip[i], clip[i + 1], 3)
        a = ref.analyze(clip[i], clip[i + 1])
        assert len(mids) == 3 and mids[0].shape == clip[i].shape and mids[0].dtype == np.uint8
        if a.is_scene_change:
            assert all(np.array_equal(m, clip[i]) for m in mids)        # :385-388
        elif a.hud_coverage > 0.01:
            for m, t in zip(mids, (0.25, 0.5, 0.75)):                   # :399-409
                assert np.array_equal(m[ref.hud_mask], (clip[i] if t < 0.5 else clip[i + 1])[ref.hud_mask])
    want = g["rows"][:8, 6].astype(int)
    assert p.get_stats()["total"] == 8 and p.get_stats()["scene_change"] == int((want == 2).sum())
    p.reset_stats()
    assert p.get_stats() == {"total": 0}


def test_process_pair_analyses_twice(engine, clip_and_gold):
    clip, _ = clip_and_gold
    p = AdaptivePipeline()
    p.load()
    for i in range(3):
        res = p.process_pair(clip[i], clip[i + 1], 3, 1.0)
    assert len(p.router.frame_history) == 6                             # :443 + :382 -> two pushes per pair
    assert len(res.frames) == 5 and res.model_used == "AdaptivePipeline"
    assert set(res.extra_info["analysis"]) == {"motion_mean", "motion_max", "has_particles", "is_scene_change",
                                               "hud_coverage", "recommended_model"}
    assert res.extra_info["routing_stats"]["total"] == 3


def test_process_clip_single_rank_equals_serial(engine, clip_and_gold):
    clip, g = clip_and_gold
    p = AdaptivePipeline()
    p.load()
    got = {}
    out = p.process_clip(clip, 3, on_pair=lambda j, mids, a: got.__setitem__(j, a.hud_coverage))
    assert out["routes"].tolist() == g["rows"][:, 6].astype(int).tolist()
    assert [got[j] for j in range(len(g["rows"]))] == g["rows"][:, 5].tolist()
    assert out["stats"]["total"] == len(g["rows"])
    # a shard starting mid-clip reproduces the serial HUD state through the warm-up replay
    p2 = AdaptivePipeline()
    p2.load()
    got2 = {}
    out2 = p2.process_clip(clip, 3, rank=1, world=2, on_pair=lambda j, mids, a: got2.__setitem__(j, a.hud_coverage))
    lo, hi = out2["range"]
    assert (lo, hi) == (8, 16) and [got2[j] for j in range(lo, hi)] == g["rows"][lo:hi, 5].tolist()


def test_rife_model_interface(engine, oracle_model):
    m = get_available_models()["rife"]()
    m.load()
    fr = gen_clip(1, 270, 480, seed=0, cut_every=0, fast=None)
    mids = m.interpolate(fr[0], fr[1], 3)
    ref = ifnet_ref.interpolate_ref(oracle_model, fr[0], fr[1], [0.25, 0.5, 0.75])
    assert len(mids) == 3 and all(np.abs(a.astype(int) - b.astype(int)).max() <= 6 for a, b in zip(mids, ref))
    assert m.interpolate(fr[0], fr[1], 1, timestamps=[0.5])[0].shape == (270, 480, 3)
    res = m.process_pair(fr[0], fr[1], 3, 1.333)
    assert len(res.frames) == 5 and res.frames[0].shape == (359, 639, 3) and res.inference_time_ms > 0
    rec = m.interpolate_recursive(fr[0], fr[1], 2)
    assert len(rec) == 3 and np.array_equal(rec[1], m.interpolate(fr[0], fr[1], 1)[0])
    fp = RIFEModel(precision="fp32")
    fp.load()
    a = fp.interpolate(fr[0], fr[1], 3)
    assert all(np.abs(x.astype(int) - y.astype(int)).max() <= 1 for x, y in zip(a, ref))
    uhd = RIFEModel(uhd_mode=True)
    uhd.load()
    assert uhd.interpolate(fr[0], fr[1], 1)[0].shape == (270, 480, 3)
    engine.load_weights(m.state_dict())


def test_metrics_reference_kats(engine):                               # /root/reference tests/test_metrics.py
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (100, 100, 3), dtype=np.uint8)
    b = rng.integers(0, 256, (100, 100, 3), dtype=np.uint8)
    assert compute_psnr_simple(a, a) == float("inf")
    assert compute_psnr_simple(np.zeros_like(a), np.full_like(a, 255)) == 0.0
    assert compute_psnr_simple(a, b) == compute_psnr_simple(b, a) == metrics_ref.psnr_simple(a, b)   # bit-identical
    assert np.isfinite(compute_psnr_simple(a[:50, :50].copy(), b))
    assert abs(compute_ssim_simple(a, a) - 1.0) < 1e-3
    assert abs(compute_ssim_simple(a, b) - metrics_ref.ssim_simple(a, b)) < 1e-9
    assert abs(compute_ssim_simple(a, b) - compute_ssim_simple(b, a)) < 1e-3
    assert abs(compute_ssim_simple(a[:50, :50].copy(), b) - metrics_ref.ssim_simple(a[:50, :50].copy(), b)) < 1e-9
    ev = QualityEvaluator()
    r = ev.evaluate([a, b], [b, b])
    assert r.psnr == float("inf") or np.isfinite(r.psnr)
    assert abs(ev.evaluate_pair(a, b)["ssim"] - metrics_ref.ssim_simple(a, b)) < 1e-9
    with pytest.raises(ValueError):
        ev.evaluate([a], [a, b])


def test_psnr_golden_from_reference(engine, golden):
    g = golden("psnr.npz")
    n, H, W, seed = g["clip"].tolist()
    fr = gen_clip(n, H, W, seed=seed, cut_every=0, fast=None)
    pred = noisy_copy(fr, seed=5, amp=3)
    assert psnr_batch(pred, fr).tolist() == g["psnr_noise"].tolist()   # bit-identical to the reference function
    assert [compute_psnr_simple(fr[i], fr[i + 1]) for i in range(n)] == g["psnr_cross"].tolist()
    assert compute_psnr_simple(fr[0][::2, ::2].copy(), fr[0]) == float(g["psnr_resize"][0])


def test_run_model_benchmark_keys(engine):
    m = RIFEModel()
    m.load()
    fr = gen_clip(2, 128, 192, seed=1, cut_every=0, fast=None)
    tr = [{"frame0": fr[0], "frame1": fr[2], "gt": [fr[1]]}]
    out = run_model_benchmark(m, tr, num_intermediate=1, target_scale=1.0)
    for k in ("time_mean_ms", "time_p99_ms", "throughput_fps", "vram_peak_mb", "psnr_mean", "ssim_mean"):
        assert k in out


def test_clip_streamer_equals_batch_api():
    """The pipelined clip API (three streams, two slots) returns exactly what interpolate_batch returns, slot reuse
    included."""
    m = RIFEModel()
    m.load()
    ts = [0.25, 0.5, 0.75]
    clip = gen_clip(13, 96, 160, seed=11, cut_every=0, fast=None)
    want = [m.interpolate_batch(clip[3 * b:3 * b + 3], clip[3 * b + 1:3 * b + 4], ts) for b in range(4)]
    torch.cuda.synchronize()           # the engine's workspace is shared: no other stream may be mid-inference
    host = torch.from_numpy(clip).pin_memory()
    st = m.clip_streamer(pairs=3, H=96, W=160, timestamps=ts, depth=2)
    pending = []
    for b in range(4):
        pending.append((st.submit(host[3 * b:3 * b + 4]), b))
        if len(pending) == 2:                       # collect before the slot is reused
            slot, b0 = pending.pop(0)
            assert np.array_equal(st.result(slot).numpy(), want[b0])
    slot, b0 = pending.pop(0)
    assert np.array_equal(st.result(slot).numpy(), want[b0])


@pytest.mark.parametrize("rank,world,batch", [(0, 1, 8), (0, 1, 5), (1, 2, 8), (2, 3, 4)])
def test_process_clip_batched_equals_per_pair_path(engine, clip_and_gold, rank, world, batch):
    """The device-resident batched clip path (one statistics read and one IFNet call per batch, per-pair HUD masks) gives
    the per-pair path's routes, HUD coverage and FRAMES (interpolate(): cut -> copies of frame0, HUD composite) bit for bit,
    for a whole clip and for a shard that starts mid-clip and holds only its part of the frames (`base`)."""
    from vfi_b200 import sharding
    clip, g = clip_and_gold
    n = len(clip) - 1
    lo, hi = sharding.shard_range(n, rank, world)
    base = max(0, lo - 16)
    p = AdaptivePipeline(enable_vfimamba=False)
    p.load()
    dev = torch.from_numpy(clip[base:hi + 1]).cuda()
    out = torch.empty((hi - lo, 3) + clip.shape[1:], dtype=torch.uint8, device="cuda")
    res = p.process_clip_batched(dev, 3, rank, world, batch=batch, out=out, n_pairs=n, base=base)
    assert res["range"] == (lo, hi)
    assert res["routes"][:hi - lo].tolist() == g["rows"][lo:hi, 6].astype(int).tolist()
    assert res["hud_coverage"][:hi - lo].tolist() == g["rows"][lo:hi, 5].tolist()
    assert [a.is_scene_change for a in res["analyses"]] == g["rows"][lo:hi, 4].astype(bool).tolist()
    assert p.get_stats()["total"] == hi - lo
    # frames: against the per-pair path fed the same history
    q = AdaptivePipeline(enable_vfimamba=False)
    q.load()
    want = {}
    q.process_clip(clip, 3, rank=rank, world=world, on_pair=lambda j, mids, a: want.__setitem__(j, np.stack(mids)))
    got = out.cpu().numpy()
    for j in range(lo, hi):
        assert np.array_equal(got[j - lo], want[j]), j
    if world == 1:
        assert res["stats"]["scene_change"] == int((g["rows"][:, 4] == 1).sum())
        assert res["stats"]["recommended_vfimamba"] == int((g["rows"][:, 6] == 1).sum())


def test_evaluate_sharded_sums(engine):
    """Sharded PSNR/SSIM evaluator: a rank's partial sums over its range are the sums of the per-frame reference values
    (integer SSE exactly), and the ranges of all ranks tile the set."""
    from vfi_b200.evaluation import metrics as M
    fr = gen_clip(10, 180, 320, seed=21, cut_every=0, fast=None)
    pred = noisy_copy(fr, seed=5)
    pred[3] = fr[3]                                                     # one identical pair -> psnr inf, counted apart
    whole = M.evaluate_sharded(pred, fr, chunk=4)
    ps = [metrics_ref.psnr_simple(pred[i], fr[i]) for i in range(11)]
    ss = [metrics_ref.ssim_simple(pred[i], fr[i]) for i in range(11)]
    sse = sum(int(((pred[i].astype(np.int64) - fr[i]) ** 2).sum()) for i in range(11))
    assert whole["n"] == 11 and whole["n_identical"] == 1 and whole["sse_sum"] == sse and whole["psnr"] == float("inf")
    assert abs(whole["psnr_sum"] - sum(p for p in ps if np.isfinite(p))) < 1e-9
    assert abs(whole["ssim"] - np.mean(ss)) < 1e-9
    parts = [M.evaluate_range(lambda a, b: pred[a:b], lambda a, b: fr[a:b], *M_range, 3) for M_range in ((0, 4), (4, 8), (8, 11))]
    tot = np.sum(np.array(parts), axis=0)
    assert tot[2] == 11 and tot[3] == 1 and int(tot[4]) == sse and abs(tot[1] - whole["ssim_sum"]) < 1e-9


def test_clip_streamer_yuv420_output():
    """output='yuv420': every result frame is cv2's I420 conversion of the RGB result, at half the download bytes."""
    import cv2
    m = RIFEModel()
    m.load()
    ts = [0.25, 0.5, 0.75]
    clip = gen_clip(6, 96, 160, seed=12, cut_every=0, fast=None)
    host = torch.from_numpy(clip).pin_memory()
    rgb = m.clip_streamer(pairs=3, H=96, W=160, timestamps=ts)
    yuv = m.clip_streamer(pairs=3, H=96, W=160, timestamps=ts, output="yuv420")
    assert yuv.d2h_bytes * 2 == rgb.d2h_bytes
    for b in range(2):
        want = rgb.result(rgb.submit(host[3 * b:3 * b + 4])).numpy().copy()
        got = yuv.result(yuv.submit(host[3 * b:3 * b + 4])).numpy()
        assert got.shape == (3, 3, 144, 160)
        for p in range(3):
            for k in range(3):
                assert np.array_equal(got[p, k], cv2.cvtColor(want[p, k], cv2.COLOR_RGB2YUV_I420))
    with pytest.raises(ValueError):
        m.clip_streamer(pairs=1, H=95, W=160, timestamps=ts, output="yuv420")
