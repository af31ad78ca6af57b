"""Independent checks of the PCL half (VoxelGrid, StatisticalOutlierRemoval, RadiusOutlierRemoval).

PCL is absent from this image, so the C restatement in oracle/ cannot be pinned against the library itself ("parity
unpinned").  What can be done is to stop the check from being self-referential: here the oracle AND
the CUDA path are compared with restatements that share no code with either --
  * k nearest neighbours by scipy.spatial.cKDTree (an exact k-d tree search, as FLANN's single-tree exact search is),
  * voxel membership by numpy floor / unique / bincount in float32, centroids in float64,
  * mean / sample standard deviation / threshold by numpy in float64.
Call sites in the reference: PointCloudPostProcessor.cpp:20-24, 38-41 (SOR), MappingSystem.cpp:297-307 (VoxelGrid, ROR).
"""
import numpy as np
import pytest
from scipy.spatial import cKDTree


def _cloud(P, n, seed, outliers=0.01, box=(40, 12, 40)):
    rng = np.random.default_rng(seed)
    xyz = rng.random((n, 3), dtype=np.float32) * np.array(box, np.float32)
    no = int(n * outliers)
    xyz[:no] = (rng.random((no, 3), dtype=np.float32) - np.float32(1 / 6)) * (np.array(box, np.float32) * np.float32(1.5))
    pts = np.zeros(n, P)
    pts["x"], pts["y"], pts["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    pts["r"] = rng.integers(0, 256, n); pts["g"] = rng.integers(0, 256, n); pts["b"] = rng.integers(0, 256, n); pts["a"] = 255
    return pts


def _xyz(pts):
    return np.stack([pts["x"], pts["y"], pts["z"]], 1)


def np_voxel_grid(pts, leaf):
    """VoxelGrid by numpy: f32 `p * (1/leaf)` then floor (PCL's key arithmetic), keys relative to the cloud minimum,
    per-voxel mean in float64.  Returns (sorted unique keys, counts, centroids[f64], mean colours[f64])."""
    inv = np.float32(1.0) / np.float32(leaf)
    xyz = _xyz(pts)
    ijk = np.floor(xyz * inv).astype(np.int64)
    minb = np.floor(xyz.min(0) * inv).astype(np.int64)
    maxb = np.floor(xyz.max(0) * inv).astype(np.int64)
    div = maxb - minb + 1
    key = (ijk[:, 0] - minb[0]) + (ijk[:, 1] - minb[1]) * div[0] + (ijk[:, 2] - minb[2]) * div[0] * div[1]
    uk, invi, cnt = np.unique(key, return_inverse=True, return_counts=True)
    cen = np.stack([np.bincount(invi, weights=xyz[:, a].astype(np.float64)) / cnt for a in range(3)], 1)
    col = np.stack([np.bincount(invi, weights=pts[c].astype(np.float64)) / cnt for c in ("r", "g", "b")], 1)
    return uk, cnt, cen, col


def kd_mean_dist(pts, k, idx=None):
    """Mean distance to the k nearest neighbours (self excluded), cKDTree."""
    xyz = _xyz(pts).astype(np.float64)
    tree = cKDTree(xyz)
    q = xyz if idx is None else xyz[idx]
    d, _ = tree.query(q, k=k + 1, workers=-1)
    return d[:, 1:].mean(1)


def np_sor_threshold(dist, alpha):
    d = dist.astype(np.float64)
    n = len(d)
    mean = d.sum() / n
    var = ((d * d).sum() - d.sum() ** 2 / n) / (n - 1)
    return mean + alpha * np.sqrt(var)


# ---------------------------------------------------------------------------------------------------------------------
# CPU: the oracle against the independent restatements
@pytest.mark.parametrize("leaf", [0.2, 0.05])
def test_oracle_voxel_grid_vs_numpy(oracle, leaf):
    pts = _cloud(oracle.POINT_DTYPE, 60000, 1, box=(8, 3, 8))
    out, counts, pt = oracle.voxel_grid(pts, leaf)
    uk, cnt, cen, col = np_voxel_grid(pts, leaf)
    assert not pt and len(out) == len(uk)                         # same voxel set size; the oracle emits in ascending key order
    assert np.array_equal(counts, cnt)                            # per-voxel membership counts, voxel by voxel
    assert np.abs(_xyz(out).astype(np.float64) - cen).max() <= 1e-5 * leaf + 1e-6
    got_col = np.stack([out[c] for c in ("r", "g", "b")], 1).astype(np.float64)
    assert (np.abs(got_col - np.floor(col)) <= 1).all()           # colours are truncated means (+-1 LSB from f32 summation)
    assert leaf != 0.2 or cnt.max() > 1                           # the 0.2 grid really merges points


@pytest.mark.parametrize("k,alpha", [(30, 0.5), (8, 1.0), (50, 1.0)])
def test_oracle_sor_vs_ckdtree(oracle, k, alpha):
    pts = _cloud(oracle.POINT_DTYPE, 6000, 2, box=(4, 2, 4))
    kept, dist, thr, keep = oracle.sor(pts, k, alpha)
    ref = kd_mean_dist(pts, k)
    assert np.allclose(dist, ref, rtol=2e-6, atol=1e-7)           # f32 distances against f64: rounding only
    thr_ref = np_sor_threshold(dist, alpha)
    assert abs(thr - thr_ref) <= 1e-9 * thr_ref
    near = np.abs(ref - thr_ref) <= 1e-5 * thr_ref
    want = ref <= thr_ref
    assert np.array_equal(keep[~near].astype(bool), want[~near])
    assert np.array_equal(kept, pts[keep.astype(bool)])           # input order preserved
    assert 0 < keep.sum() < len(pts)


def test_oracle_radius_outlier_removal_vs_ckdtree(oracle):
    pts = _cloud(oracle.POINT_DTYPE, 8000, 3, box=(4, 2, 4))
    for radius, min_pts in ((0.25, 5), (0.5, 40)):
        kept, keep = oracle.radius_outlier_removal(pts, radius, min_pts)
        xyz = _xyz(pts).astype(np.float64)
        tree = cKDTree(xyz)
        cnt = np.array([len(v) for v in tree.query_ball_point(xyz, radius, workers=-1)])   # includes the point itself
        # points with a neighbour within rounding distance of the radius may legitimately differ
        edge = np.array([len(v) for v in tree.query_ball_point(xyz, radius * (1 + 1e-6), workers=-1)]) != \
            np.array([len(v) for v in tree.query_ball_point(xyz, radius * (1 - 1e-6), workers=-1)])
        want = cnt >= min_pts + 1
        assert np.array_equal(keep[~edge].astype(bool), want[~edge])
        assert 0 < keep.sum() < len(pts)


# ---------------------------------------------------------------------------------------------------------------------
# GPU: the CUDA path against the same restatements, at the multi-million-point scale of BASELINE.json configs[4]
@pytest.mark.gpu
def test_gpu_voxel_grid_4m_points_vs_numpy(cvr):
    n = 4_000_000
    pts = _cloud(cvr.POINT_DTYPE, n, 0)
    leaf = 0.05
    rec = cvr.Reconstruct3D(64, 64, voxel_leaf=leaf, max_points=n, filter_only=True)
    vox, passthrough = rec.VoxelGrid(pts)
    rec.close()
    uk, cnt, cen, col = np_voxel_grid(pts, leaf)
    assert not passthrough and len(vox) == len(uk)                # voxel set size exact (integer keys)
    # the output is in ascending key order, so voxel j of the GPU is voxel j of numpy: centroid and colour agree
    assert np.abs(_xyz(vox).astype(np.float64) - cen).max() <= 1e-5 * leaf + 4e-6
    got_col = np.stack([vox[c] for c in ("r", "g", "b")], 1).astype(np.float64)
    assert (np.abs(got_col - np.floor(col)) <= 1).all()
    # counts: a centroid of n members reproduces n when re-keyed only if it stays inside its voxel; check through the keys
    inv = np.float32(1.0) / np.float32(leaf)
    multi = cnt > 1
    assert multi.sum() > 1000
    # every multi-member centroid lies inside the voxel numpy assigned those members to
    xyz = _xyz(pts)
    minb = np.floor(xyz.min(0) * inv).astype(np.int64); maxb = np.floor(xyz.max(0) * inv).astype(np.int64); div = maxb - minb + 1
    ijk = np.floor(_xyz(vox[multi]) * inv).astype(np.int64) - minb
    key = ijk[:, 0] + ijk[:, 1] * div[0] + ijk[:, 2] * div[0] * div[1]
    assert (key == uk[multi]).mean() > 0.999                     # a centroid may round onto a voxel face


@pytest.mark.gpu
def test_gpu_sor_large_cloud_vs_ckdtree(cvr):
    n, k, alpha = 1_500_000, 50, 1.0
    pts = _cloud(cvr.POINT_DTYPE, n, 4)
    cfg = dict(block_size=5, num_disparities=16, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10,
               min_disparity=0, outlier_min_k=k, outlier_std_threshold=alpha)
    rec = cvr.Reconstruct3D(64, 64, config=cfg, max_points=n, filter_only=True)
    kept = rec.RemoveOutliers(pts)
    dist, st = rec.sor_debug(n)
    rec.close()
    # (1) the k-NN mean distances: 10 000 random queries against a cKDTree of the full cloud
    idx = np.random.default_rng(5).choice(n, 10000, replace=False)
    ref = kd_mean_dist(pts, k, idx)
    assert np.allclose(dist[idx], ref, rtol=2e-6, atol=1e-7)
    # (2) the statistics and the threshold from the device's own distances, recomputed by numpy in f64
    thr_ref = np_sor_threshold(dist, alpha)
    assert int(st["valid"]) == n and abs(float(st["thr"]) - thr_ref) <= 1e-9 * thr_ref
    # (3) the selection: exactly the points at or below the threshold, in input order
    keep = dist.astype(np.float64) <= float(st["thr"])
    assert int(st["kept"]) == int(keep.sum()) == len(kept)
    assert np.array_equal(kept, pts[keep])
    assert 0.5 * n < len(kept) < n


@pytest.mark.gpu
def test_gpu_radius_outlier_removal_vs_ckdtree(cvr):
    n = 200_000
    pts = _cloud(cvr.POINT_DTYPE, n, 6, box=(10, 4, 10))
    rec = cvr.Reconstruct3D(64, 64, max_points=n, filter_only=True)
    radius, min_pts = 0.3, 40
    kept = rec.RadiusOutlierRemoval(pts, radius, min_pts)
    rec.close()
    xyz = _xyz(pts).astype(np.float64)
    tree = cKDTree(xyz)
    lo = np.array([len(v) for v in tree.query_ball_point(xyz, radius * (1 - 1e-6), workers=-1)])
    hi = np.array([len(v) for v in tree.query_ball_point(xyz, radius * (1 + 1e-6), workers=-1)])
    sure_keep = lo >= min_pts + 1
    sure_drop = hi < min_pts + 1
    got = np.isin(pts.view(np.dtype((np.void, 16))), kept.view(np.dtype((np.void, 16))))
    assert got[sure_keep].all() and not got[sure_drop].any()
    assert 0 < len(kept) < n
