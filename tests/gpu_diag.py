"""Not a test: prints stage-by-stage mismatch counts of the CUDA path against the oracle (first bring-up aid).
usage: python tests/gpu_diag.py"""
import importlib
import os
import sys
import time

import os
os.environ.setdefault("CVR_KEEP_S", "1")
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402

cvr = importlib.import_module("master-thesis_b200")


def cfg(bs, D, minD=0):
    return dict(block_size=bs, num_disparities=D, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100,
                pre_filter_cap=10, min_disparity=minD, outlier_min_k=30, outlier_std_threshold=0.5)


def diag(H, W, D, bs, mode, minD=0):
    l, r = O.synth_pair(H, W, D, seed=1)
    p = O.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode, min_disparity=minD)
    ref, st = O.sgbm(l, r, p, stages=True)
    rec = cvr.Reconstruct3D(W, H, config=cfg(bs, D, minD), mode=mode)
    rec.set_profiling(True)
    t = time.time()
    got = rec.GenerateDisparityMap(l, r)
    dt = time.time() - t
    C = rec.debug_fetch(0)[:, :, :D]
    S = rec.debug_fetch(1)[:, :, :D]
    lr = rec.debug_fetch(3)

    def mm(a, b):
        d = a != b
        s = "%d/%d" % (int(d.sum()), d.size)
        if d.any():
            idx = np.argwhere(d)[0]
            s += " first@%s got %s ref %s" % (idx.tolist(), a[tuple(idx)], b[tuple(idx)])
        return s
    print("H%d W%d D%d bs%d mode%d minD%d: C %s | S %s | lr %s | final %s | %.1f ms" % (
        H, W, D, bs, mode, minD, mm(C, st["C"]), mm(S, st["S"]), mm(lr, st["raw"]), mm(got, ref), dt * 1e3), flush=True)
    print("   ", ["%s %.3f" % t for t in rec.last_timings()])
    rec.close()


if __name__ == "__main__":
    diag(48, 260, 64, 5, 0)
    diag(48, 260, 64, 27, 0)
    diag(48, 260, 64, 9, 1)
    diag(40, 200, 32, 7, 0, -16)
    diag(40, 300, 128, 5, 0)
    diag(40, 400, 192, 5, 1)
    diag(40, 450, 256, 5, 0)
    diag(375, 1242, 128, 27, 0)
