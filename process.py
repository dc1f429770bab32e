#!/usr/bin/env python
"""`process.py --vfi-model rife|adaptive` / `--method adaptive`: the CLI the reference's README.md:66-79
documents (never implemented there).  Thin shim over the registry; frames come from a synthetic clip
or a directory of PNGs."""
import argparse
import glob
import json
import os

import cv2
import numpy as np

import vfi_b200  # noqa: F401
from vfi_b200.scripts.run_benchmarks import get_available_models
from vfi_b200.synth import gen_clip


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--vfi-model', default='rife', choices=['rife', 'adaptive'])
    ap.add_argument('--method', default=None, choices=[None, 'adaptive', 'rife'])
    ap.add_argument('--input', default=None, help='directory of PNG frames (default: synthetic clip)')
    ap.add_argument('--output', default=None, help='directory for interpolated PNGs')
    ap.add_argument('--pairs', type=int, default=4)
    ap.add_argument('--num-frames', type=int, default=3)
    ap.add_argument('--random-weights', action='store_true',
                    help='run without train_log/flownet.pkl on seeded random-init IFNet weights (synthetic runs only)')
    a = ap.parse_args()
    name = a.method or a.vfi_model
    model = get_available_models()[name](**({'allow_random_weights': True} if a.random_weights else {}))
    model.load()
    if a.input:
        frames = [cv2.cvtColor(cv2.imread(f), cv2.COLOR_BGR2RGB) for f in sorted(glob.glob(os.path.join(a.input, '*.png')))]
    else:
        frames = list(gen_clip(a.pairs, 1080, 1920, seed=1))
    n = 0
    for i in range(len(frames) - 1):
        mids = model.interpolate(frames[i], frames[i + 1], a.num_frames)
        if a.output:
            os.makedirs(a.output, exist_ok=True)
            for k, m in enumerate(mids):
                cv2.imwrite(os.path.join(a.output, f'f{i:05d}_{k}.png'), cv2.cvtColor(np.ascontiguousarray(m), cv2.COLOR_RGB2BGR))
        n += len(mids)
    stats = model.get_stats() if hasattr(model, 'get_stats') else {}
    print(json.dumps({'model': model.info.name, 'interpolated_frames': n, 'routing': stats}))


if __name__ == '__main__':
    main()
