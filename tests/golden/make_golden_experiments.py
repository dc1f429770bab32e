#!/usr/bin/env python
"""Golden vectors for the experiment-runner call sites, produced by the REFERENCE's
own functions imported unmodified from /root/reference/scripts/run_experiments.py:

  calc_psnr / calc_ssim (:292-312), calc_motion (:359-362) on cv2.resize'd 480x270 grays as the adaptive branch calls it
  (:849-852), estimate_motion_complexity (:449-476, incl. cv2.Canny), and rife_interpolate (:365-386) -- the latter with
  RIFEModel.get() patched to hand back the ORACLE IFNet (the real one is a missing submodule) and Tensor.cuda() patched to
  the identity (no GPU here): padding, BGR flip, scale and crop are the reference's own lines.

    python tests/golden/make_golden_experiments.py
"""
import os
import sys

import cv2
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/scripts")
sys.path.insert(0, "/root/reference")

import vfi_b200  # noqa: E402,F401
from oracle import ifnet_ref  # noqa: E402
from vfi_b200 import ifnet_params  # noqa: E402
from vfi_b200.synth import gen_clip, noisy_copy  # noqa: E402

import run_experiments as rx  # noqa: E402  (the reference)

CLIP = dict(n_pairs=6, H=540, W=960, seed=7, cut_every=0, fast=(2, 4))
RIFE_CLIP = dict(n_pairs=1, H=100, W=150, seed=8, cut_every=0, fast=None)


def main():
    rgb = gen_clip(**CLIP)
    bgr = np.ascontiguousarray(rgb[..., ::-1])
    pred = np.ascontiguousarray(noisy_copy(rgb, seed=5, amp=4)[..., ::-1])
    n = CLIP["n_pairs"]
    psnr = [rx.calc_psnr(pred[i], bgr[i]) for i in range(n + 1)] + [rx.calc_psnr(bgr[i], bgr[i + 1]) for i in range(n)]
    ssim = [rx.calc_ssim(pred[i], bgr[i]) for i in range(n + 1)] + [rx.calc_ssim(bgr[i], bgr[i + 1]) for i in range(n)]
    ident = [rx.calc_psnr(bgr[0], bgr[0]), rx.calc_ssim(bgr[0], bgr[0])]
    gray = [cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in bgr]
    motion = [float(rx.calc_motion(cv2.resize(gray[i], (480, 270)), cv2.resize(gray[i + 1], (480, 270)))) for i in range(n)]
    cx = [float(rx.estimate_motion_complexity(gray[i], gray[i + 1])[0]) for i in range(n)]
    dens = [float(np.mean(cv2.Canny(g, 100, 200) > 0)) for g in gray]

    # rife_interpolate through the reference's own lines, oracle IFNet underneath
    net = ifnet_ref.Model()
    net.flownet.load_state_dict(ifnet_params.random_init(1234), strict=True)
    rx.RIFEModel._instance = net
    torch.Tensor.cuda = lambda self, *a, **k: self
    small = np.ascontiguousarray(gen_clip(**RIFE_CLIP)[..., ::-1])
    rife = {}
    with torch.no_grad():
        for scale in (0.5, 1.0):
            for t in (0.25, 0.5):
                rife[f"rife_s{scale}_t{t}"] = rx.rife_interpolate(small[0], small[1], t, scale)
    np.savez_compressed(os.path.join(HERE, "experiments.npz"), clip=np.array([n, CLIP["H"], CLIP["W"], CLIP["seed"]]),
                        fast=np.array(CLIP["fast"]), psnr=np.array(psnr), ssim=np.array(ssim), ident=np.array(ident),
                        motion=np.array(motion), complexity=np.array(cx), canny_density=np.array(dens),
                        rife_clip=np.array([RIFE_CLIP["n_pairs"], RIFE_CLIP["H"], RIFE_CLIP["W"], RIFE_CLIP["seed"]]), **rife)
    print("motion", motion)
    print("complexity", cx)
    print("density", dens)
    print("psnr", psnr[:3], "ssim", ssim[:3], ident)


if __name__ == "__main__":
    main()
