"""Side measurement (not a bench line): the other BASELINE.json shapes through the same library, with per-stage times.
usage: python profiles/side_shapes.py [spi|hd|hd_hh|spi_bs5] [--batch N]
  spi     1280x720  D=192 MODE_HH (8 paths) + LR check, block_size 27 (reference parameters)
  spi_bs5 the same at block_size 5 (a meaningful map)
  hd      1920x1080 D=256 MODE_SGBM, block_size 27
  hd_hh   1920x1080 D=256 MODE_HH"""
import argparse
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("shape", nargs="?", default="spi")
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--steps", type=int, default=3)
a = ap.parse_args()
cvr = importlib.import_module("master-thesis_b200")
synth = importlib.import_module("master-thesis_b200.synth")
SH = {"spi": (1280, 720, 192, 27, 1, 1, 24), "spi_bs5": (1280, 720, 192, 5, 1, 1, 24), "hd": (1920, 1080, 256, 27, 0, 0, 8),
      "hd_hh": (1920, 1080, 256, 27, 1, 0, 8), "kitti": (1242, 375, 128, 27, 0, 0, 104)}
W, H, D, bs, mode, d12, B = SH[a.shape]
B = a.batch or B
cfg = cvr.parse_config(None)
cfg.update(block_size=bs, num_disparities=D)
pairs = [synth.synth_pair(H, W, D, seed=s, color=True) for s in range(2)]
Ls = np.stack([pairs[i % 2][0] for i in range(B)])
Rs = np.stack([pairs[i % 2][1] for i in range(B)])
rec = cvr.Reconstruct3D(W, H, config=cfg, max_batch=B, mode=mode, disp12_max_diff=d12)
rec.set_profiling(True)
rec.ProcessFrames(Ls, Rs, cloud_capacity=262144)
acc = {}
t0 = time.perf_counter()
for it in range(a.steps):
    rec.ProcessFrames(Ls, Rs, cloud_capacity=262144)
    for n, ms in rec.last_timings():
        acc.setdefault(n, []).append(ms)
dt = (time.perf_counter() - t0) / a.steps
dev_ms = sum(np.mean(v) for k, v in acc.items() if k not in ("h2d", "out"))
cells = (W - D) * H * D
print("%s %dx%d D=%d bs=%d mode=%s batch=%d: %.1f frames/s on the device (%.2f ms/frame; host call incl. copies %.1f frames/s)"
      % (a.shape, W, H, D, bs, "HH" if mode else "SGBM", B, B / (dev_ms * 1e-3), dev_ms / B, B / dt))
print("  us/frame: " + " ".join("%s=%.0f" % (k, np.mean(v) * 1000 / B) for k, v in acc.items()))
print("  cost volume %.0f MB/frame" % (cells * 2 / 1e6))
rec.close()
