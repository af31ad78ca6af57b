"""Side measurement (not a bench line): BASELINE.json configs[4] -- 20 M-point cloud, voxel leaf 0.05, SOR k = 50, alpha = 1.0.
usage: python profiles/side_filter_stress.py [--points 20000000]"""
import argparse
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--points", type=int, default=20_000_000)
a = ap.parse_args()
cvr = importlib.import_module("master-thesis_b200")
n = a.points
rng = np.random.default_rng(0)
xyz = rng.random((n, 3), dtype=np.float32) * np.array([40, 12, 40], np.float32)
nout = n // 100
xyz[:nout] = (rng.random((nout, 3), dtype=np.float32) - np.float32(1 / 6)) * np.array([60, 18, 60], np.float32)   # 1 % outliers, 1.5x box
pts = np.zeros(n, cvr.POINT_DTYPE)
pts["x"], pts["y"], pts["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
pts["r"] = rng.integers(0, 256, n); pts["a"] = 255
cfg = cvr.parse_config(None)
cfg.update(outlier_min_k=50, outlier_std_threshold=1.0)
rec = cvr.Reconstruct3D(64, 64, config=cfg, voxel_leaf=0.05, max_points=n, filter_only=True)
rec.set_profiling(True)
for it in range(2):
    t0 = time.perf_counter(); vox, passthrough = rec.VoxelGrid(pts); t1 = time.perf_counter()
    tv = rec.last_timings()
    kept = rec.RemoveOutliers(vox); t2 = time.perf_counter()
    ts = rec.last_timings()
print("points %d -> voxels %d (passthrough %d) -> kept %d" % (n, len(vox), passthrough, len(kept)))
print("voxel grid: host call %.1f ms; stages %s" % ((t1 - t0) * 1e3, ["%s %.2f" % t for t in tv]))
print("SOR k=50  : host call %.1f ms; stages %s" % ((t2 - t1) * 1e3, ["%s %.2f" % t for t in ts]))
rec.close()
