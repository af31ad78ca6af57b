"""GPU: BASELINE.json's full-size configurations through size-independent properties (the oracle would take minutes
there): determinism, batch == single frame, structural invariants of the disparity map, voxel/SOR invariants on a
multi-million-point cloud.  The small-size bit-exactness against cv2 / the oracle is in test_gpu_parity.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cfg(bs, D, k=30, alpha=0.5):
    return dict(block_size=bs, num_disparities=D, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10,
                min_disparity=0, outlier_min_k=k, outlier_std_threshold=alpha)


def _disparity_invariants(disp, D, H, W, min_valid):
    assert disp.shape == (H, W) and disp.dtype == np.int16
    assert (disp[:, :D] == -16).all()                           # columns x < maxDisparity are never computed (A.0)
    v = disp[disp != -16]
    assert v.size > min_valid * H * (W - D)
    assert v.min() >= 0 and v.max() <= (D - 1) * 16             # x16 fixed point within the search range


@pytest.mark.parametrize("H,W,D,bs,mode", [(1080, 1920, 256, 27, 0), (720, 1280, 192, 27, 1), (720, 1280, 192, 5, 1)])
def test_full_size_configs(cvr, oracle, H, W, D, bs, mode):
    """configs[2] (StereoPi-shape 1280x720 D=192, 8-path MODE_HH with LR check) and configs[3] (1920x1080 D=256)."""
    l, r = oracle.synth_pair(H, W, D, seed=5)
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D), mode=mode)
    a = rec.GenerateDisparityMap(l, r)
    b = rec.GenerateDisparityMap(l, r)
    assert np.array_equal(a, b)                                 # deterministic
    # at block_size 27 with 8 paths the aggregated cost saturates and uniqueness rejects almost everything
    _disparity_invariants(a, D, H, W, 0.5 if (bs == 5 or mode == 0) else 0.001)
    if bs == 5:                                                 # ground truth of the generator: most pixels within one disparity
        yy = np.arange(H)[:, None] * np.ones((1, W))
        gt = np.floor(0.15 * D + 0.5 * D * yy / H)
        ok = a != -16
        err = np.abs(a[ok] / 16.0 - gt[ok])
        assert np.median(err) <= 1.0
    rec.close()
    # the generic (global-memory state) path kernels must agree with the shared-memory ones at a size where only they fit
    import os
    os.environ["CVR_PATH_KERNELS"] = "generic"
    try:
        rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D), mode=mode)
        assert np.array_equal(rec.GenerateDisparityMap(l, r), a)
        rec.close()
    finally:
        del os.environ["CVR_PATH_KERNELS"]


def test_kitti_batch_equals_single_and_golden(cvr, oracle):
    """configs[1]: KITTI shape, frames in a batch give the same disparity as frame-by-frame calls and as cv2."""
    from conftest import load_golden
    g = load_golden("kitti_d128_bs27")
    H, W, D, bs = 375, 1242, 128, 27
    pairs = [oracle.synth_pair(H, W, D, seed=s, color=True) for s in (0, 3, 4)]
    Ls = np.stack([p[0] for p in pairs]); Rs = np.stack([p[1] for p in pairs])
    rec = cvr.Reconstruct3D(W, H, max_batch=3)
    disp, clouds = rec.ProcessFrames(Ls, Rs, want_disp=True, cloud_capacity=200000)
    assert np.array_equal(disp[0], g["disp"])                   # seed 0 is the golden KITTI frame (cv2)
    for i in range(3):
        assert np.array_equal(rec.GenerateDisparityMap(Ls[i], Rs[i]), disp[i])
        c = clouds[i]
        assert 100 < len(c) < 200000 and np.isfinite(c["x"]).all() and (c["z"] > 0).all()
    rec.close()


def test_filter_stress_cloud(cvr, oracle):
    """configs[4] (scaled to 4 M points so the GPU suite stays short; same generator: uniform cloud + 1 % far outliers,
    leaf 0.05, SOR k = 50, alpha = 1.0): invariants of the voxel grid and of the outlier removal."""
    n = 4_000_000
    rng = np.random.default_rng(0)
    xyz = rng.random((n, 3), dtype=np.float32) * np.array([40, 12, 40], np.float32)
    nout = n // 100
    # outliers in a 1.5x larger box: 1200 x 360 x 1200 voxels still fit PCL's int32 index (a 3x box would overflow it and
    # make VoxelGrid pass the cloud through, which test_gpu_parity covers)
    xyz[:nout] = (rng.random((nout, 3), dtype=np.float32) - np.float32(1 / 6)) * np.array([60, 18, 60], np.float32)
    pts = np.zeros(n, cvr.POINT_DTYPE)
    pts["x"], pts["y"], pts["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    pts["r"] = rng.integers(0, 256, n); pts["g"] = 7; pts["b"] = 9; pts["a"] = 255
    rec = cvr.Reconstruct3D(64, 64, config=_cfg(5, 16, k=50, alpha=1.0), voxel_leaf=0.05, max_points=n, filter_only=True)
    vox, passthrough = rec.VoxelGrid(pts)
    assert not passthrough
    # every voxel centroid lies inside its voxel; voxel keys are strictly increasing (sorted, unique)
    inv = np.float32(1.0) / np.float32(0.05)
    mn = pts[["x", "y", "z"]].copy()
    mnx, mny, mnz = pts["x"].min(), pts["y"].min(), pts["z"].min()
    mxx, mxy = pts["x"].max(), pts["y"].max()
    b0 = np.floor(np.array([mnx, mny, mnz], np.float32) * inv).astype(np.int64)
    dv = np.floor(np.array([mxx, mxy], np.float32) * inv).astype(np.int64) - b0[:2] + 1
    ijk = np.stack([np.floor(vox[f] * inv).astype(np.int64) for f in ("x", "y", "z")], 1) - b0
    key = ijk[:, 0] + ijk[:, 1] * dv[0] + ijk[:, 2] * dv[0] * dv[1]
    # a centroid can round onto a voxel face, so allow equality with the neighbour but require global order
    assert (np.diff(key) >= -1).mean() > 0.999
    # the number of voxels equals the number of distinct input keys (exact: integer keys)
    pk = [np.floor(pts[f] * inv).astype(np.int64) for f in ("x", "y", "z")]
    pkey = (pk[0] - b0[0]) + (pk[1] - b0[1]) * dv[0] + (pk[2] - b0[2]) * dv[0] * dv[1]
    assert len(vox) == len(np.unique(pkey))
    kept = rec.RemoveOutliers(vox[:1_500_000])
    assert 0.5 * 1_500_000 < len(kept) <= 1_500_000
    # idempotence-like property: the kept set is a subsequence of the input (order preserved)
    kb = kept.tobytes(); assert len(kb) == len(kept) * 16
    idx = np.flatnonzero(np.isin(vox[:1_500_000].view(np.dtype((np.void, 16))), kept.view(np.dtype((np.void, 16)))))
    assert len(idx) >= len(kept)
    rec.close()
