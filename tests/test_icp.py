"""point-to-point ICP with PCL's stop rules (MappingSystem.cpp:321-341: correspondence distance 5,
50 iterations, transformation epsilon 1e-8, fitness epsilon 1).  PCL is absent => parity unpinned; the C restatement in
oracle/ is cross-checked here against an independent numpy / scipy.cKDTree restatement (CPU), and the CUDA path against the
oracle (GPU).  Tolerances: transform entries 1e-5 (rotation) / 1e-5 x scene size (translation), aligned points 1e-4 m,
iteration count and stop reason equal."""
import numpy as np
import pytest
from scipy.spatial import cKDTree


def _clouds(P, n, seed, angle=0.03, shift=(0.05, -0.02, 0.08), noise=0.0, extra=0):
    rng = np.random.default_rng(seed)
    xyz = rng.random((n, 3)).astype(np.float32) * np.array([10, 4, 10], np.float32)
    tgt = np.zeros(n, P)
    tgt["x"], tgt["y"], tgt["z"] = xyz.T
    c, s = np.cos(angle), np.sin(angle)
    R = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], np.float64)
    sx = (xyz[: n - extra].astype(np.float64) @ R.T + np.array(shift)) + noise * rng.standard_normal((n - extra, 3))
    src = np.zeros(n - extra, P)
    src["x"], src["y"], src["z"] = sx.astype(np.float32).T
    src["r"] = 200
    return src, tgt


def np_icp(src, tgt, max_dist=5.0, max_iter=50, trans_eps=1e-8, fit_eps=1.0):
    """Independent restatement: cKDTree neighbours, numpy SVD (Kabsch / Umeyama without scaling), PCL's stop rules."""
    S = np.stack([src["x"], src["y"], src["z"]], 1).astype(np.float64)
    Tg = np.stack([tgt["x"], tgt["y"], tgt["z"]], 1).astype(np.float64)
    tree = cKDTree(Tg)
    T = np.eye(4)
    prev, it, state = np.finfo(np.float64).max, 0, 0
    while True:
        d, j = tree.query(S)
        ok = d <= max_dist
        if ok.sum() < 3:
            break
        a, b = S[ok], Tg[j[ok]]
        ma, mb = a.mean(0), b.mean(0)
        H = (b - mb).T @ (a - ma) / len(a)
        U, _, Vt = np.linalg.svd(H)
        D = np.diag([1, 1, np.sign(np.linalg.det(U) * np.linalg.det(Vt))])
        R = U @ D @ Vt
        t = mb - R @ ma
        Ti = np.eye(4); Ti[:3, :3] = R; Ti[:3, 3] = t
        S = S @ R.T + t
        T = Ti @ T
        it += 1
        if it >= max_iter:
            state = 1; break
        if 0.5 * (np.trace(R) - 1) >= 1 - trans_eps and t @ t <= trans_eps:
            state = 2; break
        mse = float((d[ok] ** 2).mean())
        if abs(mse - prev) < 1e-12:
            state = 3; break
        if abs(mse - prev) / prev < fit_eps:
            state = 4; break
        prev = mse
    return T, S, it, state


@pytest.mark.parametrize("kw", [dict(), dict(fit_eps=1e-9), dict(fit_eps=1e-9, max_iter=3), dict(max_dist=0.3, fit_eps=1e-6)])
def test_oracle_icp_vs_numpy_ckdtree(oracle, kw):
    src, tgt = _clouds(oracle.POINT_DTYPE, 1500, 1, noise=0.002, extra=100)
    T, aligned, it, fit, state, nc = oracle.icp(src, tgt, **kw)
    Tn, Sn, itn, staten = np_icp(src, tgt, **kw)
    assert (it, state) == (itn, staten)
    assert np.abs(T[:3, :3] - Tn[:3, :3]).max() < 1e-5 and np.abs(T[:3, 3] - Tn[:3, 3]).max() < 1e-4
    assert np.abs(np.stack([aligned["x"], aligned["y"], aligned["z"]], 1) - Sn).max() < 1e-4
    assert np.array_equal(aligned["r"], src["r"])                  # colours ride along untouched
    if not kw:
        assert it == 2 and state == 4      # the reference's fitness epsilon of 1 ends the alignment after two iterations


def test_oracle_icp_degenerate(oracle):
    src, tgt = _clouds(oracle.POINT_DTYPE, 50, 2)
    far = src.copy(); far["x"] += 1000.0
    T, aligned, it, fit, state, nc = oracle.icp(far, tgt)           # nothing within 5 m: PCL reports "not enough correspondences"
    assert state == 0 and it == 0 and nc == 0 and np.array_equal(T, np.eye(4, dtype=np.float32))
    T, aligned, it, fit, state, nc = oracle.icp(tgt, tgt, fit_eps=1e-9)   # already aligned: identity, stops on the transformation epsilon
    assert state == 2 and np.allclose(T, np.eye(4), atol=1e-6) and fit < 1e-10
    sums = np.zeros(17); sums[0] = 4                                 # coplanar correspondences: a rotation is still returned (det +1)
    a = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0]], float); b = a @ np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1.0]]).T + [2, 0, 1]
    sums[1:4] = a.sum(0); sums[4:7] = b.sum(0); sums[7:16] = (b.T @ a).ravel()
    Tu = oracle.umeyama_from_sums(sums)
    assert abs(np.linalg.det(Tu[:3, :3].astype(np.float64)) - 1) < 1e-5 and np.allclose(a @ Tu[:3, :3].T + Tu[:3, 3], b, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("kw,okw", [(dict(), dict()), (dict(euclidean_fitness_epsilon=1e-9), dict(fit_eps=1e-9)),
                                    (dict(max_correspondence_distance=0.3, euclidean_fitness_epsilon=1e-6), dict(max_dist=0.3, fit_eps=1e-6))])
def test_gpu_icp_vs_oracle(cvr, oracle, kw, okw):
    src, tgt = _clouds(cvr.POINT_DTYPE, 3000, 3, noise=0.002, extra=200)
    rec = cvr.Reconstruct3D(64, 64, max_points=4096, filter_only=True)
    T, aligned, res = rec.AlignICP(src, tgt, **kw)
    To, ao, ito, fito, so, nco = oracle.icp(src, tgt, **okw)
    assert (res.iterations, res.state, res.correspondences) == (ito, so, nco)
    assert np.abs(T[:3, :3] - To[:3, :3]).max() < 1e-5 and np.abs(T[:3, 3] - To[:3, 3]).max() < 1e-4
    for f in ("x", "y", "z"):
        assert np.abs(aligned[f] - ao[f]).max() < 1e-4
    assert abs(res.fitness - fito) <= 1e-5 * max(fito, 1e-9) + 1e-12
    assert np.array_equal(aligned["r"], src["r"])
    # source points outside the target's bounding box and a cloud with no correspondences
    far = src.copy(); far["x"] += 1000.0
    T2, _, r2 = rec.AlignICP(far, tgt)
    assert r2.state == 0 and r2.converged == 0 and np.array_equal(T2, np.eye(4, dtype=np.float32))
    off = src.copy(); off["x"] += 3.0
    T3, a3, r3 = rec.AlignICP(off, tgt)
    T3o, a3o, it3, f3, s3, n3 = oracle.icp(off, tgt)
    assert (r3.iterations, r3.state, r3.correspondences) == (it3, s3, n3) and np.abs(T3 - T3o).max() < 1e-4
    rec.close()


@pytest.mark.gpu
def test_gpu_icp_large_clouds_vs_ckdtree(cvr):
    """200 k points: the grid search against cKDTree through the result (the oracle's brute force would take minutes)."""
    src, tgt = _clouds(cvr.POINT_DTYPE, 200_000, 4, angle=0.01, shift=(0.03, 0.01, -0.02), noise=0.001)
    rec = cvr.Reconstruct3D(64, 64, max_points=200_000, filter_only=True)
    T, aligned, res = rec.AlignICP(src, tgt, euclidean_fitness_epsilon=1e-9)
    Tn, Sn, itn, staten = np_icp(src, tgt, fit_eps=1e-9)
    assert (res.iterations, res.state) == (itn, staten)
    assert np.abs(T[:3, :3] - Tn[:3, :3]).max() < 1e-5 and np.abs(T[:3, 3] - Tn[:3, 3]).max() < 2e-4
    assert np.abs(np.stack([aligned["x"], aligned["y"], aligned["z"]], 1) - Sn).max() < 5e-4
    rec.close()
