"""N>1 host path on CPU: world_size-2 (and 3) gloo.  Each rank takes its contiguous pair range, replays the
HUD pushes it needs from before its range (sharding.run_sharded), analyses its pairs with the ORACLE router
standing in for the kernels, and the counters are summed by the one all-reduce.  The merged per-pair results
must equal the serial reference run (tests/golden/router.npz) bit for bit -- including HUD coverage, which is
the only state that crosses shard boundaries."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import vfi_b200  # noqa: F401
    from oracle import router_ref
    from vfi_b200 import sharding
    from vfi_b200.synth import gen_clip
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = np.load(os.path.join(ROOT, "tests", "golden", "router.npz"))
    n, H, W, seed = g["clip"].tolist()
    clip = gen_clip(n, H, W, seed=seed, cut_every=int(g["cut_every"]), fast=tuple(g["fast"].tolist()))
    r = router_ref.RouterRef()
    code = {"rife": 0, "vfimamba": 1, "scene_change": 2}

    def process(j):
        a = r.analyze(clip[j], clip[j + 1])
        return [a.motion_mean, a.motion_max, a.motion_std, float(a.has_particles), float(a.is_scene_change),
                a.hud_coverage, code[a.recommended_model], a.confidence]

    (lo, hi), rows = sharding.run_sharded(
        n, rank, world, is_cut=lambda j: router_ref.scene_ssim(clip[j], clip[j + 1]) < 0.65,
        push_hud=lambda j: r.detect_hud(clip[j]), process=process)
    rows = np.array(rows).reshape(-1, 8)
    cnt = torch.tensor([len(rows), int((rows[:, 6] == 0).sum()), int((rows[:, 6] == 1).sum()),
                        int((rows[:, 6] == 2).sum())], dtype=torch.int64)
    cnt = sharding.reduce_counters(cnt)
    flags = sharding.gather_flags(torch.tensor(rows[:, 6], dtype=torch.int8), n, rank, world)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), lo=lo, hi=hi, rows=rows, cnt=cnt.numpy(), flags=flags.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_router_equals_serial(tmp_path, world):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    g = np.load(os.path.join(ROOT, "tests", "golden", "router.npz"))
    parts = [np.load(tmp_path / f"r{r}.npz") for r in range(world)]
    assert int(parts[0]["lo"]) == 0 and int(parts[-1]["hi"]) == len(g["rows"])
    merged = np.concatenate([p["rows"] for p in parts])
    assert np.array_equal(merged, g["rows"])                       # HUD coverage included: warm-up replay is exact
    want = [len(g["rows"]), int((g["rows"][:, 6] == 0).sum()), int((g["rows"][:, 6] == 1).sum()),
            int((g["rows"][:, 6] == 2).sum())]
    for p in parts:
        assert p["cnt"].tolist() == want                           # all-reduce result on every rank
        assert p["flags"].tolist() == g["rows"][:, 6].astype(int).tolist()


def test_single_process_paths_need_no_group():
    sys.path.insert(0, ROOT)
    import vfi_b200  # noqa: F401
    from vfi_b200 import sharding
    v = torch.tensor([1, 2, 3])
    assert sharding.reduce_counters(v.clone()).tolist() == [1, 2, 3]
    assert sharding.gather_flags(torch.tensor([0, 2, 1], dtype=torch.int8), 3, 0, 1).tolist() == [0, 2, 1]


def _sum_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import vfi_b200  # noqa: F401
    from vfi_b200 import sharding
    from vfi_b200.evaluation import metrics as M
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_range(10, rank, world)
    # a rank's evaluate_range() vector: [sum psnr, sum ssim, n, n_inf, sum sse] over its frames (synthetic per-frame values)
    psnr = [30.0 + i for i in range(lo, hi)]
    part = [sum(psnr), sum(0.9 + 0.001 * i for i in range(lo, hi)), float(hi - lo), float(rank == 0), float(sum(1000 * i for i in range(lo, hi)))]
    res = M.reduce_sums(part, (lo, hi), None, "cpu")
    cov = sharding.gather_flags(torch.tensor([0.5 * i for i in range(lo, hi)], dtype=torch.float64), 10, rank, world)
    np.savez(os.path.join(out_dir, f"s{rank}.npz"), n=res["n"], sse=res["sse_sum"], psnr_sum=res["psnr_sum"], ssim=res["ssim"],
             inf=res["n_identical"], cov=cov.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_metric_sums_all_reduce(tmp_path, world):
    """Config 5's one collective (evaluation.metrics.reduce_sums) and the float64 flag gather over gloo: every rank ends
    with the whole-set sums (integer entries exact) and the clip-order coverage vector."""
    mp.spawn(_sum_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        p = np.load(tmp_path / f"s{r}.npz")
        assert int(p["n"]) == 10 and int(p["sse"]) == sum(1000 * i for i in range(10)) and int(p["inf"]) == 1
        assert abs(float(p["psnr_sum"]) - sum(30.0 + i for i in range(10))) < 1e-9
        assert abs(float(p["ssim"]) - np.mean([0.9 + 0.001 * i for i in range(10)])) < 1e-12
        assert p["cov"].tolist() == [0.5 * i for i in range(10)]
