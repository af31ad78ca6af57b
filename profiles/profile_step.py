"""One warm-up step + `--steps` steps of the bench workload (KITTI shape, batch per GPU as in bench.py), nothing else:
the command ncu wraps.  usage: python profiles/profile_step.py [--batch 36] [--steps 1]"""
import argparse
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=36)
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--mode", type=int, default=0)
a = ap.parse_args()
cvr = importlib.import_module("master-thesis_b200")
synth = importlib.import_module("master-thesis_b200.synth")
W, H, D = 1242, 375, 128
pairs = [synth.synth_pair(H, W, D, seed=s, color=True) for s in range(4)]
Ls = np.stack([pairs[i % 4][0] for i in range(a.batch)])
Rs = np.stack([pairs[i % 4][1] for i in range(a.batch)])
rec = cvr.Reconstruct3D(W, H, max_batch=a.batch, mode=a.mode)
rec.set_profiling(True)
for it in range(1 + a.steps):
    rec.ProcessFrames(Ls, Rs, cloud_capacity=65536)
    print("step", it, "launches so far", rec.launch_count(), ["%s %.3f" % t for t in rec.last_timings()], flush=True)
rec.close()
