"""Randomised parity sweep (tool, not collected by pytest): CUDA disparity vs the CPU oracle over random shapes and SGBM
parameters, including batches and the int16-saturation regime.  usage: python tests/gpu_fuzz.py [--n 60] [--seed 1]"""
import argparse
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=60)
ap.add_argument("--seed", type=int, default=1)
a = ap.parse_args()
cvr = importlib.import_module("master-thesis_b200")
from oracle import oracle as O  # noqa: E402

O.build()
rng = np.random.default_rng(a.seed)
bad = 0
t0 = time.time()
for it in range(a.n):
    D = int(rng.choice([16, 32, 48, 64, 96, 128, 160, 192, 256]))
    bs = int(rng.choice([1, 3, 5, 7, 9, 11, 15, 21, 27]))
    mode = int(rng.integers(0, 2))
    minD = int(rng.choice([0, 0, 0, -16, 16, -5, 7]))
    H = int(rng.integers(8, 90))
    W = int(D + abs(minD) + (rng.integers(1, 400) if rng.integers(0, 4) else rng.integers(400, 1400)))
    B = int(rng.choice([1, 1, 2, 3, 5]))
    uniq = int(rng.choice([0, 5, 10, 15]))
    cap = int(rng.choice([10, 31, 63]))
    spw = int(rng.choice([0, 50, 100]))
    d12 = int(rng.choice([-1, 0, 1, 2]))
    kind = int(rng.integers(0, 3))
    pairs = []
    for i in range(B):
        seed = int(rng.integers(0, 1 << 30))
        if kind == 0:
            l, r = O.synth_pair(H, W, D, seed, color=False)
        elif kind == 1:
            l, r = O.satdrift_pair(H, W, D, seed)
        else:
            g = np.random.default_rng(seed)
            l = g.integers(0, 256, (H, W), dtype=np.uint8); r = g.integers(0, 256, (H, W), dtype=np.uint8)
        pairs.append((np.ascontiguousarray(l), np.ascontiguousarray(r)))
    cfg = dict(block_size=bs, num_disparities=D, uniqueness_ratio=uniq, speckle_range=2, speckle_window_size=spw, pre_filter_cap=cap,
               min_disparity=minD, outlier_min_k=8, outlier_std_threshold=0.5)
    p = O.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode, min_disparity=minD, uniqueness_ratio=uniq,
                                speckle_window_size=spw, pre_filter_cap=cap, disp12_max_diff=d12)
    try:
        rec = cvr.Reconstruct3D(W, H, config=cfg, mode=mode, disp12_max_diff=d12, max_batch=B)
    except cvr.CvrError as e:
        print("skip (rejected):", D, bs, mode, minD, str(e)[:60]); continue
    Ls = np.stack([q[0] for q in pairs]); Rs = np.stack([q[1] for q in pairs])
    disp, _ = rec.ProcessFrames(Ls, Rs, stages=1, want_disp=True)
    for i in range(B):
        ref = O.sgbm(pairs[i][0], pairs[i][1], p)
        if not np.array_equal(disp[i], ref):
            bad += 1
            print("MISMATCH", dict(D=D, bs=bs, mode=mode, minD=minD, H=H, W=W, B=B, uniq=uniq, cap=cap, spw=spw, d12=d12, kind=kind, frame=i),
                  int((disp[i] != ref).sum()))
    rec.close()
print("fuzz: %d configurations, %d mismatching frames, %.0f s" % (a.n, bad, time.time() - t0))
sys.exit(1 if bad else 0)
