#!/usr/bin/env python
"""Generates the golden fixtures in this directory.  Run HERE (the build container), where the
reference tree is mounted at /root/reference; the fixtures travel to the GPU box, the reference does not.

    python tests/golden/make_golden.py

What comes from the REFERENCE ITSELF (imported unmodified from /root/reference):
  * psnr.npz          evaluation/metrics.py:299-315 compute_psnr_simple on seeded frame pairs
  * base_helpers.npz  models/base.py:175-229 to_tensor / pad_to_multiple / unpad / to_numpy
  * router.npz        models/novel/adaptive_pipeline.py:74-286 AdaptiveRouter.analyze over a synthetic clip
                      (cuts, a fast segment, a static HUD), every ContentAnalysis field + the HUD mask bits.
                      skimage is not installed here, so `skimage.metrics.structural_similarity` is served
                      by the oracle restatement (oracle/metrics_ref.py) through a sys.modules shim: the SSIM
                      *value* is therefore oracle-defined (PARITY UNPINNED), everything else in analyze()
                      (cv2 gray/resize/Farneback/Laplacian/morphology, numpy reductions, the decision tree)
                      is the reference's own code.
What comes from the ORACLE (the IFNet source is a missing third-party submodule):
  * ifnet_small.npz   oracle/ifnet_ref.py on a 96x160 pair, seeded random-init weights.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)

import vfi_b200  # noqa: E402,F401
from vfi_b200 import ifnet_params  # noqa: E402
from vfi_b200.synth import gen_clip, noisy_copy  # noqa: E402
from oracle import ifnet_ref, metrics_ref  # noqa: E402

ROUTER_CLIP = dict(n_pairs=16, H=360, W=640, seed=2, cut_every=6, fast=(8, 10))


def _import_reference():
    shim = types.ModuleType("skimage")
    shim_m = types.ModuleType("skimage.metrics")
    shim_m.structural_similarity = lambda a, b, **kw: metrics_ref.ssim_gray_u8(a, b)
    shim.metrics = shim_m
    sys.modules.setdefault("skimage", shim)
    sys.modules.setdefault("skimage.metrics", shim_m)
    sys.path.insert(0, REF)
    import importlib
    ref_metrics = importlib.import_module("evaluation.metrics")
    ref_base = importlib.import_module("models.base")
    ref_adapt = importlib.import_module("models.novel.adaptive_pipeline")
    return ref_metrics, ref_base, ref_adapt


def main():
    ref_metrics, ref_base, ref_adapt = _import_reference()

    # ---- PSNR (reference function) ---------------------------------------------------------
    fr = gen_clip(5, 120, 200, seed=11, cut_every=0, fast=None)
    pred = noisy_copy(fr, seed=5, amp=3)
    ps = [ref_metrics.compute_psnr_simple(pred[i], fr[i]) for i in range(len(fr))]
    ps_cross = [ref_metrics.compute_psnr_simple(fr[i], fr[i + 1]) for i in range(len(fr) - 1)]
    small = fr[0][::2, ::2].copy()
    ps_resize = ref_metrics.compute_psnr_simple(small, fr[0])
    np.savez_compressed(os.path.join(HERE, "psnr.npz"), clip=np.array([5, 120, 200, 11]), psnr_noise=np.array(ps),
                        psnr_cross=np.array(ps_cross), psnr_resize=np.array([ps_resize]))

    # ---- base.py helpers (reference class) -------------------------------------------------
    class _M(ref_base.BaseModel):
        info = None
        def load(self): pass
        def interpolate(self, *a, **k): return []
        def upscale(self, f, scale=1.0): return f
    m = _M(device="cpu")
    f = gen_clip(0, 50, 70, seed=12, cut_every=0, fast=None)[0]
    t = m.to_tensor(f)
    tp, pad = m.pad_to_multiple(t, 32)
    back = m.to_numpy(m.unpad(tp, f.shape[0], f.shape[1]))
    trunc = m.to_numpy((t * 0 + 0.99999).clone())      # 0.99999*255 = 254.99 -> 254: truncation, not rounding
    np.savez_compressed(os.path.join(HERE, "base_helpers.npz"), frame=f, tensor=t.numpy(), padded=tp.numpy(),
                        pad=np.array(pad), back=back, trunc=trunc, ts3=np.array(m.get_default_timestamps(3)),
                        ts1=np.array(m.get_default_timestamps(1)))

    # ---- router (reference class; SSIM served by the oracle shim) ---------------------------
    clip = gen_clip(**ROUTER_CLIP)
    r = ref_adapt.AdaptiveRouter()
    r2 = ref_adapt.AdaptiveRouter(motion_threshold_low=1.5, motion_threshold_high=10.0, particle_threshold=0.055)
    rows, rows2, masks = [], [], []
    code = {"rife": 0, "vfimamba": 1, "scene_change": 2}
    for i in range(len(clip) - 1):
        a = r.analyze(clip[i], clip[i + 1])
        rows.append([a.motion_mean, a.motion_max, a.motion_std, float(a.has_particles), float(a.is_scene_change),
                     a.hud_coverage, code[a.recommended_model], a.confidence])
        b = r2.analyze(clip[i], clip[i + 1])   # tighter thresholds: exercises the vfimamba / 0.7 / particle branches
        rows2.append([b.motion_mean, b.motion_max, b.motion_std, float(b.has_particles), float(b.is_scene_change),
                      b.hud_coverage, code[b.recommended_model], b.confidence])
        masks.append(np.packbits(r.hud_mask if (r.hud_mask is not None and not a.is_scene_change and len(r.frame_history) >= 5)
                                 else np.zeros(clip[i].shape[:2], bool)))
    # the pieces on their own (reference methods)
    lap = [float(ref_adapt.cv2.Laplacian(ref_adapt.cv2.cvtColor(c, ref_adapt.cv2.COLOR_RGB2GRAY),
                                         ref_adapt.cv2.CV_64F).var()) for c in clip]
    motion = [r.compute_motion(clip[i], clip[i + 1])[:3] for i in (0, 1, 9)]
    np.savez_compressed(os.path.join(HERE, "router.npz"), clip=np.array(list(ROUTER_CLIP.values())[:4]),
                        cut_every=ROUTER_CLIP["cut_every"], fast=np.array(ROUTER_CLIP["fast"]),
                        rows=np.array(rows, dtype=np.float64), rows_tight=np.array(rows2, dtype=np.float64),
                        tight=np.array([1.5, 10.0, 0.055]), masks=np.stack(masks), laplacian_var=np.array(lap),
                        motion_pairs=np.array([0, 1, 9]), motion=np.array(motion, dtype=np.float64))

    # ---- IFNet (oracle) --------------------------------------------------------------------
    sd = ifnet_params.random_init(1234)
    net = ifnet_ref.Model()
    net.flownet.load_state_dict(sd, strict=True)
    fr = gen_clip(1, 96, 160, seed=0, cut_every=0, fast=None)
    ts = [0.25, 0.5, 0.75]
    u8, f32 = ifnet_ref.interpolate_ref(net, fr[0], fr[1], ts, return_float=True)
    np.savez_compressed(os.path.join(HERE, "ifnet_small.npz"), clip=np.array([1, 96, 160, 0]), ts=np.array(ts),
                        weight_seed=1234, out_u8=np.stack(u8), out_f32=np.stack(f32).astype(np.float32))
    for n in sorted(os.listdir(HERE)):
        if n.endswith(".npz"):
            print(n, os.path.getsize(os.path.join(HERE, n)))


if __name__ == "__main__":
    main()
