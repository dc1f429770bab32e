"""Row N4: the reference's on-disk clip formats (triplet clips, interval directories) written and read back, the triplet
index against the reference's own create_triplet_index arithmetic, and the prefetching reader's ordering."""
import json

import cv2
import numpy as np
import pytest

import vfi_b200  # noqa: F401
from vfi_b200 import data_io
from vfi_b200.synth import gen_clip


def test_triplet_index_matches_reference_docstring():
    """preprocess_video.py:351-362: 30 -> 120 fps: triplet 0 = input[0,1] -> gt[1,2,3]; triplet 1 = input[1,2] -> gt[5,6,7]."""
    ins = [f"frame_{i:06d}.png" for i in range(3)]
    gts = [f"frame_{i:06d}.png" for i in range(9)]
    tr = data_io.create_triplet_index(ins, gts, 30, 120)
    assert [t["gt_intermediate_indices"] for t in tr] == [[1, 2, 3], [5, 6, 7]]
    assert tr[0]["temporal_positions"] == [0.25, 0.5, 0.75] and tr[1]["num_intermediate"] == 3
    assert tr[1]["input_frame_0"] == ins[1] and tr[1]["gt_frame_end"] == gts[8] and tr[1]["triplet_id"] == 1
    assert data_io.create_triplet_index(ins, gts[:8], 30, 120)[-1]["triplet_id"] == 0     # gt_end out of range -> skipped
    with pytest.raises(ValueError):
        data_io.create_triplet_index([], gts, 30, 120)


def test_clip_and_interval_round_trip(tmp_path):
    fr = gen_clip(8, 72, 128, seed=3, cut_every=0, fast=None)               # 9 frames = 3 input frames at 4x
    data_io.write_clip(tmp_path / "clipA", fr, temporal_scale=4, spatial_scale=1.0)
    cd = data_io.load_clip_data(tmp_path / "clipA")
    assert set(cd) == {"metadata", "triplets", "input_dir", "gt_dir"} and len(cd["triplets"]) == 2
    assert cd["metadata"]["spatial_scale"] == 1.0 and cd["metadata"]["num_triplets"] == 2
    t0 = cd["triplets"][0]
    assert np.array_equal(data_io.load_image(cd["input_dir"] / t0["input_frame_1"]), fr[4])     # PNG is lossless, RGB order kept
    assert np.array_equal(data_io.load_image(cd["gt_dir"] / t0["gt_intermediate"][1]), fr[2])
    with pytest.raises(FileNotFoundError, match="Metadata not found"):
        data_io.load_clip_data(tmp_path / "nope")
    with pytest.raises(FileNotFoundError, match="Could not load image"):
        data_io.load_image(tmp_path / "missing.png")
    bgr = np.ascontiguousarray(fr[..., ::-1])
    data_io.write_interval(tmp_path, "clipA", "interval_0000", bgr, {"mean_flow": 1.5})
    iv = data_io.load_interval(tmp_path, "clipA", "interval_0000")
    assert len(iv["keyframes"]) == 5 and len(iv["midpoints"]) == 4 and iv["motion_stats"] == {"mean_flow": 1.5}
    assert np.array_equal(iv["keyframes"][1], bgr[2]) and np.array_equal(iv["midpoints"][0], bgr[1])
    assert iv["meta"]["keyframe_count"] == 5 and iv["meta"]["resolution"] == "128x72"
    assert data_io.load_interval(tmp_path, "clipA", "interval_0007") is None
    assert [m["interval_id"] for m in data_io.list_available_intervals(tmp_path, "clipA")] == ["interval_0000"]
    assert data_io.list_available_intervals(tmp_path, "other") == []


def test_prefetch_reader_order_and_reuse(tmp_path):
    fr = gen_clip(19, 48, 64, seed=5, cut_every=0, fast=None)
    paths = []
    for i, f in enumerate(fr):
        p = tmp_path / f"f_{i:03d}.png"
        cv2.imwrite(str(p), cv2.cvtColor(f, cv2.COLOR_RGB2BGR))
        paths.append(p)
    got = [(i, t.numpy().copy()) for i, t in data_io.PrefetchReader(paths, threads=4, ahead=3, pin=False)]
    assert [i for i, _ in got] == list(range(20)) and all(np.array_equal(a, fr[i]) for i, a in got)
    bgr = [t.numpy().copy() for _, t in data_io.PrefetchReader(paths[:2], threads=1, ahead=1, rgb=False, pin=False)]
    assert np.array_equal(bgr[1], fr[1][..., ::-1])
    with pytest.raises(FileNotFoundError):
        list(data_io.PrefetchReader([tmp_path / "missing.png"], pin=False))


@pytest.mark.gpu
def test_run_clip_benchmark_from_disk(engine, tmp_path):
    """run_clip_benchmark over a clip directory = the reference's run_model_benchmark result shape, and the numbers of the
    in-memory loop on the same frames."""
    from vfi_b200.models.sota.rife_wrapper import RIFEModel
    from vfi_b200.scripts.run_benchmarks import run_clip_benchmark, run_model_benchmark
    fr = gen_clip(12, 128, 192, seed=9, cut_every=0, fast=None)
    data_io.write_clip(tmp_path / "c", fr, temporal_scale=4, spatial_scale=1.0)
    m = RIFEModel()
    m.load()
    out = run_clip_benchmark(m, data_io.load_clip_data(tmp_path / "c"), save_outputs=True, output_dir=tmp_path / "out")
    assert out["num_triplets"] == 3 and set(out) == {"model", "model_info", "num_triplets", "quality_summary", "speed_summary",
                                                     "quality_per_triplet", "speed_per_triplet"}
    assert set(out["speed_summary"]["time_ms"]) == {"mean", "std", "p50", "p95", "p99"} and out["speed_summary"]["throughput_fps"] > 0
    mem = run_model_benchmark(m, [{"frame0": fr[4 * i], "frame1": fr[4 * i + 4], "gt": list(fr[4 * i + 1:4 * i + 4])} for i in range(3)], 3, 1.0)
    assert abs(out["quality_summary"]["psnr"]["mean"] - mem["psnr_mean"]) < 1e-9
    assert abs(out["quality_summary"]["ssim"]["mean"] - mem["ssim_mean"]) < 1e-9
    saved = cv2.cvtColor(cv2.imread(str(tmp_path / "out" / "triplet_0001" / "frame_00.png")), cv2.COLOR_BGR2RGB)
    assert np.array_equal(saved, fr[4])
    assert run_clip_benchmark(m, {**data_io.load_clip_data(tmp_path / "c"), "triplets": []}) == {"error": "No successful evaluations"}
