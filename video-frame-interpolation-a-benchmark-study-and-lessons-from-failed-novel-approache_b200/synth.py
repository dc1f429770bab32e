"""Synthetic clip generator shared by tests and bench.py (SURVEY.md §8(d) input spec).

Per scene: a smooth low-frequency background (separable sines, one small matmul per
channel) and a fine band-limited texture, both panning by an integer velocity so that
Farneback locks on (sub-threshold motion -> 'rife' route); a window of coarse texture
that jumps 36 px/frame inside the `fast` segment (max flow > 25 -> 'vfimamba'
recommendation); a static HUD rectangle; a hard cut (new scene) every `cut_every`
frames (scene-cut SSIM << 0.65).  uint8 RGB out.
"""
import cv2
import numpy as np

N_SINES = 6


def _tex(rng, H, W, sigma, amp):
    n = rng.standard_normal((H, W)).astype(np.float32)
    n = cv2.GaussianBlur(n, (0, 0), sigma)
    return n * (amp / float(n.std()))


def _scene(rng, H, W):
    sc = {}
    sc["A"] = rng.uniform(10, 35, (3, N_SINES)).astype(np.float32)
    sc["fx"] = rng.uniform(-1 / 64, 1 / 64, (3, N_SINES)).astype(np.float32)
    sc["fy"] = rng.uniform(-1 / 64, 1 / 64, (3, N_SINES)).astype(np.float32)
    sc["ph"] = rng.uniform(0, 2 * np.pi, (3, N_SINES)).astype(np.float32)
    sc["fine"] = _tex(rng, H, W, 2.5, 4.0)
    sc["coarse"] = _tex(rng, H, W, 6.0, 30.0)
    sc["vel"] = rng.integers(-2, 3, 2)
    if not sc["vel"].any():
        sc["vel"][0] = 1
    wy, wx = int(H * rng.uniform(0.35, 0.5)), int(W * rng.uniform(0.3, 0.5))
    win = np.zeros((H, W), np.float32)
    win[wy:wy + max(16, H // 4), wx:wx + max(16, W // 5)] = 1
    sc["win"] = cv2.GaussianBlur(win, (0, 0), max(2.0, H / 90.0))
    return sc


def _scene_start(i, cut_every):
    return (i // cut_every) * cut_every if cut_every else 0


def _fast_steps(lo, hi, fast):
    """Number of frame indices j in (lo, hi] with fast[0] < j <= fast[1]."""
    if not fast:
        return 0
    return max(0, min(hi, fast[1]) - max(lo, fast[0]))


def _render(sc, pan, fshift, x, y, hud_pat, out):
    H, W = out.shape[:2]
    fine = np.roll(sc["fine"], (int(pan[1]), int(pan[0])), (0, 1))
    coarse = np.roll(sc["coarse"], (0, fshift), (0, 1))
    for c in range(3):
        ax = 2 * np.pi * sc["fx"][c][:, None] * (x[None, :] - pan[0])
        ay = 2 * np.pi * sc["fy"][c][:, None] * (y[None, :] - pan[1]) + sc["ph"][c][:, None]
        L = np.concatenate([sc["A"][c][:, None] * np.sin(ay), sc["A"][c][:, None] * np.cos(ay)], 0)
        R = np.concatenate([np.cos(ax), np.sin(ax)], 0)
        img = 128.0 + L.T @ R + fine
        img = img * (1 - sc["win"]) + (128.0 + coarse) * sc["win"]
        out[:, :, c] = np.clip(np.rint(img), 0, 255).astype(np.uint8)
    if hud_pat is not None:
        out[:hud_pat.shape[0], :hud_pat.shape[1]] = hud_pat


def gen_clip(n_pairs, H=1080, W=1920, seed=0, cut_every=60, fast=(120, 150), hud=True, start=0, threads=None, out=None):
    """Return uint8 RGB frames (n_pairs + 1, H, W, 3) = clip frames start .. start + n_pairs; pair j of the returned
    array is clip pair start + j.  Every frame is a pure function of its clip index (scene = the last multiple of
    `cut_every`, pan = velocity x frames since the scene began, fast-window shift = 36 px per frame inside `fast`), so a
    rank can generate just its own shard, and frames render on `threads` host threads (default: all cores, <= 16)."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    x = np.arange(W, dtype=np.float32)
    y = np.arange(H, dtype=np.float32)
    hud_h, hud_w = max(8, int(H * 0.08)), max(8, int(W * 0.12))
    hy, hx = np.mgrid[0:hud_h, 0:hud_w]
    hud_pat = np.stack([40 + 170 * (((hx // 6) + (hy // 6)) % 2), 200 - 120 * ((hx // 9) % 2),
                        60 + 100 * ((hy // 5) % 2)], -1).astype(np.uint8) if hud else None
    frames = out if out is not None else np.empty((n_pairs + 1, H, W, 3), np.uint8)
    idx = list(range(start, start + n_pairs + 1))
    scenes = {}
    for s0 in sorted({_scene_start(i, cut_every) for i in idx}):
        scenes[s0] = _scene(np.random.default_rng(seed * 1000003 + s0 + 17), H, W)

    def one(k):
        i = idx[k]
        s0 = _scene_start(i, cut_every)
        sc = scenes[s0]
        pan = sc["vel"].astype(np.int64) * (i - s0)
        _render(sc, pan, 36 * _fast_steps(s0, i, fast), x, y, hud_pat, frames[k])

    nt = threads if threads is not None else min(16, os.cpu_count() or 1)
    if nt > 1 and len(idx) > 1:
        with ThreadPoolExecutor(nt) as ex:
            list(ex.map(one, range(len(idx))))
    else:
        for k in range(len(idx)):
            one(k)
    return frames


def noisy_copy(frames, seed=5, amp=3):
    """Config-5 style prediction: gt + U{-amp..amp} noise, clipped."""
    rng = np.random.default_rng(seed)
    n = rng.integers(-amp, amp + 1, frames.shape, dtype=np.int16)
    return np.clip(frames.astype(np.int16) + n, 0, 255).astype(np.uint8)
