"""Diagnostic: vfi_ssim_gray_u8 / vfi_psnr_ssim_u8 against the oracle per size and per frame."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import vfi_b200  # noqa
from vfi_b200.engine import Engine
from vfi_b200.synth import gen_clip, noisy_copy
from oracle import metrics_ref

e = Engine.get(0)
for (H, W, n) in [(1080, 1920, 3), (1080, 1920, 1), (360, 640, 3), (45, 71, 3), (129, 250, 3), (1440, 2560, 2), (600, 129, 3), (1080, 250, 3), (300, 1920, 3)]:
    fr = gen_clip(n - 1, H, W, seed=3, cut_every=0, fast=None, hud=H > 100) if n > 1 else gen_clip(1, H, W, seed=3, cut_every=0, fast=None)[:1]
    pred = noisy_copy(fr, seed=6)
    got = e.ssim(pred, fr, 1).cpu().numpy()
    got2 = e.psnr_ssim(pred, fr)[2].cpu().numpy()
    want = np.array([metrics_ref.ssim_simple(pred[i], fr[i]) for i in range(len(fr))])
    print(H, W, n, "ssim err", np.abs(got - want), "fused err", np.abs(got2 - want), flush=True)
