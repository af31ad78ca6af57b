"""CPU: the oracle (oracle/cvr_oracle.c) against the cv2-generated golden fixtures (tests/golden/make_golden.py).
This is what pins the OpenCV half of the oracle: every fixture was produced by the real library the reference
calls (cv2 4.13.0 StereoSGBM / cvtColor / normalize / meanStdDev / threshold / reprojectImageTo3D)."""
import zlib

import numpy as np
import pytest

from conftest import GOLDEN_BIG, GOLDEN_LARGE, GOLDEN_SMALL, load_golden


def _inputs(O, g):
    H, W, D, bs, mode, minD, seed = [int(v) for v in g["meta"]]
    lb, rb = O.synth_pair(H, W, D, seed, color=True)
    assert zlib.crc32(lb.tobytes()) == int(g["crc"][0]) and zlib.crc32(rb.tobytes()) == int(g["crc"][1]), "synthetic generator drifted"
    p = O.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode, min_disparity=minD)
    return lb, rb, p


@pytest.mark.parametrize("name", GOLDEN_SMALL + GOLDEN_LARGE[:2])
def test_sgbm_oracle_matches_cv2(oracle, name):
    g = load_golden(name)
    lb, rb, p = _inputs(oracle, g)
    gl, gr = oracle.gray_bgr(lb), oracle.gray_bgr(rb)
    assert zlib.crc32(gl.tobytes()) == int(g["gray_crc"][0]) and zlib.crc32(gr.tobytes()) == int(g["gray_crc"][1])
    if "gray_left" in g:
        assert np.array_equal(gl, g["gray_left"])
    disp, st = oracle.sgbm(gl, gr, p, stages=True)
    assert np.array_equal(disp, g["disp"]), "disparity differs from cv2 in %d pixels" % int((disp != g["disp"]).sum())
    # order-independence guard of the saturating S accumulation: all path costs non-negative
    assert st["minL"] >= 0
    p2 = dict(p, speckleWindowSize=0)
    assert np.array_equal(oracle.sgbm(gl, gr, p2), g["disp_nospeckle"])


@pytest.mark.parametrize("name", GOLDEN_SMALL + GOLDEN_LARGE[:2])
def test_prune_and_reproject_oracle_matches_cv2(oracle, name):
    g = load_golden(name)
    pruned, dispf, mask, u8, st = oracle.prune(g["disp"])
    assert np.array_equal(pruned, g["pruned"])
    assert np.array_equal(mask, g["mask"])
    assert st[2] == float(g["sigma"])
    assert np.array_equal(dispf, g["pruned"].astype(np.float32) / np.float32(16))
    f = np.float32((np.float32(959.791) + np.float32(903.7596)) / np.float32(2))
    Q = np.eye(4, dtype=np.float32)
    Q[2, 2] = np.float32(0.005 * float(f))
    xyz = oracle.reproject(g["pruned"], Q, True)
    assert zlib.crc32(xyz.tobytes()) == int(g["xyz_crc"][0])           # bitwise
    assert int((xyz[..., 2] != 10000.0).sum()) == int(g["xyz_valid_count"][0])
    if "xyz" in g:
        assert np.array_equal(xyz.view(np.uint32), g["xyz"].view(np.uint32))


@pytest.mark.parametrize("name", GOLDEN_BIG)
def test_oracle_matches_cv2_at_full_size(oracle, name):
    """BASELINE.json configs[2] / configs[3] at the reference's parameters: 1920x1080 D=256 and 1280x720 D=192 MODE_HH."""
    g = load_golden(name)
    lb, rb, p = _inputs(oracle, g)
    disp = oracle.sgbm(oracle.gray_bgr(lb), oracle.gray_bgr(rb), p)
    assert zlib.crc32(disp.tobytes()) == int(g["disp_crc"][0])
    assert np.array_equal(disp[::16], g["rows16"])
    pruned, dispf, mask, _, st = oracle.prune(disp)
    assert zlib.crc32(pruned.tobytes()) == int(g["pruned_crc"][0]) and zlib.crc32(mask.tobytes()) == int(g["mask_crc"][0])
    assert st[2] == float(g["sigma"])


def test_known_answers(oracle):
    """Known-answer micro tests of SURVEY Appendix C.2."""
    assert np.int16(np.int32(69984).astype(np.int16)) == 4448            # s16(P2) at block_size 27
    # median at borders: replicate
    a = np.array([[5, 1, 9], [7, 3, 2], [8, 6, 4]], np.int16)
    m = oracle.median3(a)
    assert m[1, 1] == 5 and m[0, 0] == sorted([5, 5, 1, 5, 5, 1, 7, 7, 3])[4]
    # speckle: component of exactly maxSize is removed, maxSize + 1 is kept
    img = np.full((12, 40), -16, np.int16)
    img[2, 1:11] = 100      # 10 pixels
    img[6, 1:12] = 100      # 11 pixels
    out = oracle.filter_speckles(img, -16, 10, 32)
    assert (out[2] == -16).all() and (out[6, 1:12] == 100).all()
    # adjacency is pairwise, not against a seed: a slow ramp stays one component
    ramp = np.full((3, 30), -16, np.int16)
    ramp[1, :] = np.arange(30) * 30
    assert np.array_equal(oracle.filter_speckles(ramp, -16, 10, 32), ramp)
    # reproject: d == min -> Z = 10000
    d = np.array([[0, 16], [32, 0]], np.int16)
    xyz = oracle.reproject(d, np.eye(4, dtype=np.float32), True)
    assert xyz[0, 0, 2] == 10000.0 and xyz[1, 1, 2] == 10000.0 and xyz[0, 1, 2] == 16.0


def test_saturation_regime_sticks_and_drifts(oracle):
    """Appendix C.2: with block_size 27 the vertical recurrence must saturate (stick at 32767) and then drift; the
    oracle follows the saturating recurrences (not the closed form) and still equals cv2 bit for bit."""
    g = load_golden("satdrift_d32_bs27")
    l, r = oracle.satdrift_pair()
    assert zlib.crc32(l.tobytes()) == int(g["crc"][0]) and zlib.crc32(r.tobytes()) == int(g["crc"][1])
    for mode in (0, 1):
        p = oracle.reference_sgbm_params(block_size=27, num_disparities=32, mode=mode)
        disp, st = oracle.sgbm(l, r, p, stages=True)
        assert st["C"].max() == 32767
        # closed form would be P2s + sum(pix); the recurrence must differ somewhere below the saturated band
        assert np.array_equal(disp, g["disp_mode%d" % mode])
        p5 = oracle.reference_sgbm_params(block_size=27, num_disparities=32, mode=mode, uniqueness_ratio=0, speckle_window_size=0)
        p5["P1"], p5["P2"] = 8, 32
        assert np.array_equal(oracle.sgbm(l, r, p5), g["disp_smallP_mode%d" % mode])


def test_pcl_side_hand_cases(oracle):
    """PCL-side restatement on hand-computable clouds (Appendix C.4)."""
    P = oracle.POINT_DTYPE
    # 8 lattice voxels, two points each (+-eps) -> 8 centroids at the lattice points, counts 2
    base = np.array([[x, y, z] for x in (0.1, 0.3) for y in (0.1, 0.3) for z in (0.1, 0.3)], np.float32)
    pts = np.zeros(16, P)
    eps = np.float32(0.01)
    for i, b in enumerate(base):
        for j, s in enumerate((-eps, eps)):
            q = pts[2 * i + j]
            q["x"], q["y"], q["z"] = b[0] + s, b[1], b[2]
            q["r"], q["g"], q["b"], q["a"] = 10 * i, 20, 30 + j, 255
    out, counts, pt = oracle.voxel_grid(pts, 0.2)
    assert len(out) == 8 and (counts == 2).all() and not pt
    assert np.allclose(np.sort(out["x"]), np.sort(base[:, 0]), atol=1e-6)
    assert set(out["b"]) == {30}                       # (30 + 31) / 2 = 30.5 -> truncated
    # int32 overflow -> passthrough
    far = pts.copy()
    far["x"][0] = 1e9
    o2, _, pt2 = oracle.voxel_grid(far, 0.2)
    assert pt2 and len(o2) == len(far)
    # SOR: one far outlier is removed for any alpha at k >= 1; all-equal distances -> sigma 0 -> keep all
    rng = np.random.default_rng(0)
    c = np.zeros(200, P)
    xyz = rng.random((200, 3)).astype(np.float32)
    c["x"], c["y"], c["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    c["x"][7] = 50.0
    kept, dist, thr, keep = oracle.sor(c, 8, 1.0)
    assert keep[7] == 0 and keep.sum() >= 190
    line = np.zeros(50, P)
    ring = np.linspace(0, 2 * np.pi, 50, endpoint=False)
    line["x"], line["y"] = np.cos(ring).astype(np.float32), np.sin(ring).astype(np.float32)
    kept2, dist2, thr2, keep2 = oracle.sor(line, 2, 0.0)
    assert keep2.sum() >= 25                           # symmetric ring: distances equal up to rounding
    # duplicates are legal
    dup = np.concatenate([c[:20], c[:20]])
    k3, d3, _, _ = oracle.sor(dup, 4, 1.0)
    assert np.isfinite(d3).all()


BM_GOLDEN = ["bm_60x300_d64_w9", "bm_61x300_d64_w21", "bm_64x260_d32_w11_minneg", "bm_64x260_d48_w15_nouniq", "bm_64x260_d48_w15_speckle",
             "bm_kitti_d128_w27"]


def bm_golden_inputs(O, g):
    H, W, D, ws, seed = [int(v) for v in g["meta"]]
    lb, rb = O.synth_pair(H, W, D, seed, color=True)
    assert zlib.crc32(lb.tobytes()) == int(g["crc"][0])
    p = {str(k): int(v) for k, v in zip(g["param_names"], g["params"])}
    return lb, rb, p


@pytest.mark.parametrize("name", BM_GOLDEN)
def test_stereobm_oracle_matches_cv2(oracle, name):
    """the StereoBM branch of the block-matcher switch, pinned against cv2.StereoBM."""
    g = load_golden(name)
    lb, rb, p = bm_golden_inputs(oracle, g)
    got = oracle.stereobm(oracle.gray_bgr(lb), oracle.gray_bgr(rb), p)
    assert np.array_equal(got, g["disp"]), "%d pixels differ from cv2.StereoBM" % int((got != g["disp"]).sum())


# ---- stereo rectification of the inputs (Reconstruct3D::RectifyImages): oracle vs cv2 fixtures ----
RECT_GOLDEN = ["rect_small_k5", "rect_small_k8_border", "rect_small_nodist_shift", "rect_small_k12", "rect_kitti_k5"]


def rect_image(H, W, ch, seed):
    img = np.random.default_rng(seed).integers(0, 256, (H, W, ch), dtype=np.uint8)
    return img[:, :, 0].copy() if ch == 1 else img


@pytest.mark.parametrize("name", RECT_GOLDEN)
def test_oracle_rectification_matches_cv2(oracle, name):
    g = load_golden(name)
    H, W, ch, seed = (int(v) for v in g["meta"])
    img = rect_image(H, W, ch, seed)
    assert zlib.crc32(img.tobytes()) == int(g["crc"][0]), "input generator drifted"
    m1, m2 = oracle.init_undistort_rectify_map(g["K"], g["D"], g["R"], g["P"], W, H)
    assert zlib.crc32(m1.tobytes()) == int(g["crc"][1]) and zlib.crc32(m2.tobytes()) == int(g["crc"][2])
    out = oracle.remap_bilinear(img, m1, m2)
    assert zlib.crc32(out.tobytes()) == int(g["crc"][3])
    if "map1" in g.files:
        assert np.array_equal(m1, g["map1"]) and np.array_equal(m2, g["map2"]) and np.array_equal(out, g["out"])
