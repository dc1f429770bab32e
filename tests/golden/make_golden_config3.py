#!/usr/bin/env python
"""Golden for BASELINE.json config 3: the REFERENCE's AdaptiveRouter.analyze
(/root/reference/models/novel/adaptive_pipeline.py:237-286, imported unmodified; skimage served by the oracle shim as
in make_golden.py) run serially over the 300-pair synthetic 1080p clip (seed 2, a cut every 60 pairs, fast segment
120-150, static HUD).  Stores per pair: route code, cut flag, HUD coverage (float64), motion mean/max/std, particle
flag, confidence, and the CRC32 of the packed HUD mask -- small arrays only; the clip is regenerated from its seed.

    python tests/golden/make_golden_config3.py         (about 5 minutes on 8 cores; run HERE, the reference is mounted)
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import _import_reference  # noqa: E402
from vfi_b200.synth import gen_clip  # noqa: E402

CLIP = dict(n_pairs=300, H=1080, W=1920, seed=2, cut_every=60, fast=(120, 150))


def main():
    _, _, ref_adapt = _import_reference()
    n = CLIP["n_pairs"]
    r = ref_adapt.AdaptiveRouter()
    code = {"rife": 0, "vfimamba": 1, "scene_change": 2}
    rows, crc = [], []
    chunk = 30
    for s in range(0, n, chunk):
        m = min(chunk, n - s)
        clip = gen_clip(m, CLIP["H"], CLIP["W"], seed=CLIP["seed"], cut_every=CLIP["cut_every"], fast=CLIP["fast"], start=s)
        for i in range(m):
            a = r.analyze(clip[i], clip[i + 1])
            rows.append([a.motion_mean, a.motion_max, a.motion_std, float(a.has_particles), float(a.is_scene_change),
                         a.hud_coverage, code[a.recommended_model], a.confidence])
            have = r.hud_mask is not None and not a.is_scene_change and len(r.frame_history) >= 5
            crc.append(zlib.crc32(np.packbits(r.hud_mask).tobytes()) if have else 0)
        print(s + m, "pairs", flush=True)
    rows = np.array(rows, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "config3_1080p.npz"), clip=np.array([n, CLIP["H"], CLIP["W"], CLIP["seed"]]),
                        cut_every=CLIP["cut_every"], fast=np.array(CLIP["fast"]), rows=rows,
                        routes=rows[:, 6].astype(np.int8), cuts=rows[:, 4].astype(np.int8), hud_coverage=rows[:, 5],
                        mask_crc32=np.array(crc, dtype=np.uint32))
    print("routes", np.bincount(rows[:, 6].astype(int), minlength=3), "cuts", int(rows[:, 4].sum()))


if __name__ == "__main__":
    main()
