"""Not a test: one small-shape pass over every kernel family, meant to run under compute-sanitizer
(profiles/sanitize.sh: memcheck and racecheck, for each path-kernel variant).  Results are also compared with the oracle so
that a sanitizer-clean run that computes nonsense does not pass unnoticed.
usage: compute-sanitizer --tool memcheck python tests/gpu_sanitize_pass.py"""
import importlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402

cvr = importlib.import_module("master-thesis_b200")


def cfg(bs, D, **kw):
    c = dict(block_size=bs, num_disparities=D, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10,
             min_disparity=0, outlier_min_k=8, outlier_std_threshold=0.5)
    c.update(kw)
    return c


def main():
    O.build()
    bad = 0
    # SGBM, both modes, a padded disparity count (HASPAD) and the no-neighbour regime (bs 27), batch of 3 through the pipelined path
    for (H, W, D, bs, mode) in ((40, 300, 64, 27, 0), (36, 280, 48, 5, 1), (24, 330, 128, 9, 0), (20, 420, 192, 5, 1), (18, 460, 256, 3, 0)):
        B = 3
        pairs = [O.synth_pair(H, W, D, seed=40 + i, color=True) for i in range(B)]
        L = np.ascontiguousarray(np.stack([p[0] for p in pairs])); R = np.ascontiguousarray(np.stack([p[1] for p in pairs]))
        rec = cvr.Reconstruct3D(W, H, config=cfg(bs, D), mode=mode, max_batch=B, voxel_leaf=0.05)
        disp, clouds = rec.ProcessFrames(L, R, want_disp=True)
        for i in range(B):
            ref = O.sgbm(O.gray_bgr(L[i]), O.gray_bgr(R[i]), O.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode))
            bad += int((disp[i] != ref).sum())
        io = cvr.Batch()
        io.mem, io.channels, io.stages = cvr.MEM_HOST, 3, 15
        io.left, io.right = L.ctypes.data, R.ctypes.data
        io.left_step = io.right_step = W * 3
        io.left_frame_stride = io.right_frame_stride = W * H * 3
        packed = np.zeros(W * H * B, cvr.POINT_DTYPE); off = np.zeros(B + 1, np.int32)
        io.cloud_packed, io.cloud_packed_capacity, io.cloud_offsets = packed.ctypes.data, len(packed), off.ctypes.data
        rec.submit_batch_raw(io, B); rec.submit_batch_raw(io, B); rec.wait_batch(); rec.wait_batch()
        for i in range(B):
            bad += int(not np.array_equal(packed[off[i]:off[i + 1]], clouds[i]))
        pts, foff, fr = rec.GatherClouds(capacity=W * H * B)
        bad += int(len(pts) != off[B])
        print("sgbm %dx%d D%d bs%d mode%d: mismatches so far %d, cloud sizes %s" % (W, H, D, bs, mode, bad, [len(c) for c in clouds]), flush=True)
        rec.close()
    # StereoBM, rectification, the stand-alone stage calls, radius filter, transform
    H, W, D = 40, 300, 64
    l, r = O.synth_pair(H, W, D, seed=7, color=True)
    c = cfg(9, D, block_matcher=cvr.MATCHER_BM, window_size=9, bm_num_disparities=D)
    rec = cvr.Reconstruct3D(W, H, config=c)
    d = rec.GenerateDisparityMap(l, r)
    bad += int((d != O.stereobm(O.gray_bgr(l), O.gray_bgr(r), O.reference_bm_params(9, D))).sum())
    rec.close()
    rec = cvr.Reconstruct3D(W, H, config=cfg(9, D), voxel_leaf=0.05)
    K = np.array([[300.0, 0, 150.0], [0, 300.0, 20.0], [0, 0, 1]]); P = np.array([[295.0, 0, 149.0, 0], [0, 295.0, 21.0, 0], [0, 0, 1, 0]])
    rec.SetRectification(K, [-0.05, 0.01, 0.0003, -0.0002, 0.0], np.eye(3), P, K, [-0.04, 0.012, -0.0001, 0.0002, 0.0], np.eye(3), P)
    rec.RectifyImages(l, r)
    disp = rec.GenerateDisparityMap(l, r)
    pruned, dispf, mask = rec.SetupFrame(disp)
    rp = O.prune(disp)
    bad += int((pruned != rp[0]).sum()) + int((mask != rp[2]).sum())
    cloud = rec.Triangulate3D(dispf, l, mask)
    rec.GeneratePointCloud(pruned, l)
    vox, _ = rec.VoxelGrid(cloud)
    kept = rec.RemoveOutliers(vox)
    rec.RadiusOutlierRemoval(vox, 0.5, 3)
    rec.TransformPointCloud(vox, np.eye(4, dtype=np.float32))
    rec.FilterCloud(cloud)
    bad += int(len(vox) != len(O.voxel_grid(cloud, 0.05)[0]))
    print("stages: voxels %d kept %d, total mismatches %d" % (len(vox), len(kept), bad), flush=True)
    rec.close()
    print("sanitize pass", "OK" if bad == 0 else "MISMATCH")
    return 0 if bad == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
