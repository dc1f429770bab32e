"""CPU oracle for the experiment-runner call sites.

TEST INFRASTRUCTURE ONLY (tests/, bench.py's cpu_baseline).  Restates /root/reference/scripts/run_experiments.py:
calc_psnr :292-302, calc_ssim :305-312, calc_motion :359-362, rife_interpolate :365-386 (on the oracle IFNet),
estimate_motion_complexity :449-476 and the adaptive branch :848-874.  cv2 (the reference's own third-party dependency)
is called directly for resize / Farneback / Canny / addWeighted / copyMakeBorder.  Pinned against the reference's own
functions by tests/golden/experiments.npz (tests/golden/make_golden_experiments.py).
"""
import cv2
import numpy as np
import torch


def calc_psnr(img1, img2):
    mse = np.mean((img1.astype(float) - img2.astype(float)) ** 2)
    if mse == 0:
        return float('inf')
    return 20 * np.log10(255.0 / np.sqrt(mse))


def calc_ssim(img1, img2):
    C1, C2 = (0.01 * 255) ** 2, (0.03 * 255) ** 2
    img1, img2 = img1.astype(float), img2.astype(float)
    mu1, mu2 = img1.mean(), img2.mean()
    s1 = ((img1 - mu1) ** 2).mean()
    s2 = ((img2 - mu2) ** 2).mean()
    s12 = ((img1 - mu1) * (img2 - mu2)).mean()
    v = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 ** 2 + mu2 ** 2 + C1) * (s1 + s2 + C2))
    return max(0.0, min(1.0, v))


def calc_motion(prev_gray, curr_gray):
    flow = cv2.calcOpticalFlowFarneback(prev_gray, curr_gray, None, 0.5, 1, 15, 1, 5, 1.1, 0)
    return float(np.mean(np.sqrt(flow[..., 0] ** 2 + flow[..., 1] ** 2)))


def thumbnails(prev_bgr, curr_bgr):
    return tuple(cv2.resize(cv2.cvtColor(f, cv2.COLOR_BGR2GRAY), (480, 270)) for f in (prev_bgr, curr_bgr))


def rife_interpolate(model, prev, curr, t, scale=0.5):
    """:365-386 with `model` = oracle.ifnet_ref.Model (CPU)."""
    h, w = prev.shape[:2]
    ph, pw = ((h - 1) // 128 + 1) * 128, ((w - 1) // 128 + 1) * 128
    p0 = cv2.copyMakeBorder(prev, 0, ph - h, 0, pw - w, cv2.BORDER_REFLECT)
    p1 = cv2.copyMakeBorder(curr, 0, ph - h, 0, pw - w, cv2.BORDER_REFLECT)
    i0 = torch.from_numpy(p0[:, :, ::-1].astype(np.float32) / 255).permute(2, 0, 1).unsqueeze(0)
    i1 = torch.from_numpy(p1[:, :, ::-1].astype(np.float32) / 255).permute(2, 0, 1).unsqueeze(0)
    with torch.no_grad():
        mid = model.inference(i0, i1, timestep=t, scale=scale)
    return (mid.squeeze(0).permute(1, 2, 0).numpy() * 255).clip(0, 255).astype(np.uint8)[:h, :w, ::-1]


def adaptive_pair(model, prev, curr, motion_thresh, rife_scale=0.5):
    """:848-874 -- (three frames, rife used, motion)."""
    g0, g1 = thumbnails(prev, curr)
    motion = calc_motion(g0, g1)
    if motion > motion_thresh:
        return [rife_interpolate(model, prev, curr, i / 4, rife_scale) for i in (1, 2, 3)], True, motion
    return [cv2.addWeighted(prev, 1 - i / 4, curr, i / 4, 0) for i in (1, 2, 3)], False, motion


def estimate_motion_complexity(prev_gray, curr_gray):
    flow = cv2.calcOpticalFlowFarneback(prev_gray, curr_gray, None, 0.5, 1, 15, 1, 5, 1.1, 0)
    mag = np.sqrt(flow[..., 0] ** 2 + flow[..., 1] ** 2)
    mean_motion = min(1.0, np.mean(mag) / 20.0)
    motion_var = min(1.0, np.std(mag) / 15.0)
    edge_density = np.mean(cv2.Canny(curr_gray, 100, 200) > 0) * 2.0
    return min(1.0, 0.5 * mean_motion + 0.3 * motion_var + 0.2 * edge_density), flow
