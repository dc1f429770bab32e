"""Host-side mirror of the reference interface: models/base.py helpers, wrapper constructors and error
behaviour, registry, RoutingStats, clip sharding.  No GPU work."""
import numpy as np
import pytest
import torch

import vfi_b200  # noqa: F401
from vfi_b200 import sharding
from vfi_b200.models.base import BaseModel, InferenceResult, ModelInfo
from vfi_b200.models.novel.adaptive_pipeline import AdaptivePipeline, AdaptiveRouter, ContentAnalysis, RoutingStats
from vfi_b200.models.sota.rife_wrapper import RIFELiteModel, RIFEModel, get_rife_model
from vfi_b200.scripts.run_benchmarks import get_available_models


class _Dummy(BaseModel):
    @property
    def info(self):
        return ModelInfo(name="dummy", type="traditional", supports_vfi=True, supports_sr=True, supports_joint=False)

    def load(self):
        self._loaded = True

    def interpolate(self, f0, f1, num_frames=3, timestamps=None):
        return [f0.copy() for _ in range(num_frames)]

    def upscale(self, frame, scale=1.333):
        return frame


def test_base_helpers_match_reference(golden):
    g = golden("base_helpers.npz")
    m = _Dummy(device="cpu")
    t = m.to_tensor(g["frame"])
    assert np.array_equal(t.numpy(), g["tensor"])
    tp, pad = m.pad_to_multiple(t, 32)
    assert np.array_equal(tp.numpy(), g["padded"]) and list(pad) == g["pad"].tolist()
    assert np.array_equal(m.to_numpy(m.unpad(tp, *g["frame"].shape[:2])), g["back"])
    assert np.array_equal(m.to_numpy(t * 0 + 0.99999), g["trunc"])
    assert m.get_default_timestamps(3) == g["ts3"].tolist() and m.get_default_timestamps(1) == g["ts1"].tolist()


def test_process_pair_contract():
    m = _Dummy(device="cpu")
    f = np.zeros((8, 8, 3), np.uint8)
    with pytest.raises(RuntimeError, match="not loaded"):
        m.process_pair(f, f)
    m.load()
    r = m.process_pair(f, f, 3, 1.0)
    assert isinstance(r, InferenceResult) and len(r.frames) == 5 and r.model_used == "dummy" and r.inference_time_ms >= 0


def test_rife_ctor_and_info(tmp_path, monkeypatch):
    m = RIFEModel()                                   # zero-arg (run_benchmarks.py:369): finds external/Practical-RIFE
    i = m.info
    assert (i.name, i.type, i.supports_vfi, i.supports_sr, i.parameters, i.requires_gpu) == \
        ("RIFE_v4.25", "sota", True, False, 9_800_000, True)
    assert RIFELiteModel().info.name == "RIFE_Lite_v4.25" and RIFELiteModel().info.parameters == 4_500_000
    assert isinstance(get_rife_model(lite=True), RIFELiteModel) and type(get_rife_model()) is RIFEModel
    with pytest.raises(FileNotFoundError, match="weights not found"):
        RIFEModel(model_dir=str(tmp_path))            # directory without train_log (rife_wrapper.py:67-72)
    with pytest.raises(RuntimeError, match="not loaded"):
        m.process_pair(np.zeros((64, 64, 3), np.uint8), np.zeros((64, 64, 3), np.uint8))
    assert m.upscale(np.zeros((30, 40, 3), np.uint8), 1.333).shape == (39, 53, 3)


def test_missing_weights_raise_unless_opted_in(monkeypatch, capsys):
    """external/Practical-RIFE/train_log holds no flownet.pkl: like the reference (rife_wrapper.py:56-72, :94-105) the
    default is an error, never silent random weights; the synthetic configs opt in explicitly."""
    monkeypatch.delenv("VFI_B200_RANDOM_WEIGHTS", raising=False)
    with pytest.raises(FileNotFoundError, match="flownet.pkl"):
        RIFEModel().state_dict()
    sd = RIFEModel(allow_random_weights=True).state_dict()
    assert "RANDOM-INIT" in capsys.readouterr().err and len(sd) == len(RIFEModel(allow_random_weights=True).state_dict())
    monkeypatch.setenv("VFI_B200_RANDOM_WEIGHTS", "1")
    assert RIFEModel().allow_random_weights


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_no_cpu_fallback():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        RIFEModel(device="cpu").load()
    from vfi_b200.evaluation.metrics import compute_psnr_simple
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        compute_psnr_simple(np.zeros((8, 8, 3), np.uint8), np.zeros((8, 8, 3), np.uint8))


def test_registry_and_adaptive_surface():
    reg = get_available_models()
    assert reg["rife"] is RIFEModel and reg["adaptive"] is AdaptivePipeline
    p = reg["adaptive"]()
    assert p.info.name == "AdaptivePipeline" and p.info.type == "novel" and p.info.parameters == 27_700_000
    r = p.router
    assert (r.motion_threshold_low, r.motion_threshold_high, r.scene_change_threshold, r.particle_threshold,
            r.hud_variance_threshold, r.hud_history_frames) == (5.0, 25.0, 0.65, 0.4, 10.0, 10)
    assert r.hud_mask is None and len(r.frame_history) == 0
    for name in ("compute_motion", "detect_scene_change", "detect_particles", "detect_hud", "analyze"):
        assert callable(getattr(AdaptiveRouter, name))
    assert p.get_stats() == {"total": 0}
    assert [f.name for f in __import__("dataclasses").fields(ContentAnalysis)] == [
        "motion_mean", "motion_max", "motion_std", "has_particles", "is_scene_change", "hud_coverage",
        "recommended_model", "confidence"]


def test_routing_stats():
    s = RoutingStats()
    for m in ("rife", "rife", "vfimamba", "scene_change"):
        s.add(m)
    d = s.to_dict()
    assert d["total"] == 4 and d["rife"] == 2 and d["rife_pct"] == 50.0 and d["scene_change_pct"] == 25.0


@pytest.mark.parametrize("n,world", [(300, 8), (300, 1), (5, 8), (16, 2), (0, 4)])
def test_shard_range_partitions(n, world):
    rs = [sharding.shard_range(n, r, world) for r in range(world)]
    assert rs[0][0] == 0 and rs[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(rs, rs[1:])) and all(lo <= hi for lo, hi in rs)
    assert max(hi - lo for lo, hi in rs) == -(-n // world)


def test_warmup_scans_back_across_cuts():
    cuts = {8, 6}
    assert sharding.warmup_start(10, lambda j: j in cuts, 1) == (4, [4, 5, 7, 9])
    assert sharding.warmup_start(10, lambda j: j in cuts, 2) == (7, [7, 9])     # process_pair pushes twice
    assert sharding.warmup_start(2, lambda j: False, 1) == (0, [0, 1])
    assert sharding.warmup_start(0, lambda j: False, 1) == (0, [])
    assert sharding.warmup_start(3, lambda j: True, 1) == (3, [])
