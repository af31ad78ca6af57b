"""GPU: the CUDA path (through the C ABI) against the CPU oracle and the cv2 golden fixtures.
Bars: disparity bit-exact; reprojected / triangulated coordinates within 1e-5 relative
(bitwise is what we assert first and report); voxel centroids within 1e-5*leaf, rgb +-1, identical voxel set and
counts; SOR kept set identical except points within 1e-5*thr of the threshold."""
import zlib

import numpy as np
import pytest

from conftest import GOLDEN_BIG, GOLDEN_LARGE, GOLDEN_SMALL, load_golden

pytestmark = pytest.mark.gpu


def _cfg(bs, D, minD=0, uniq=10, spw=100, spr=2, cap=10, k=30, alpha=0.5):
    return dict(block_size=bs, num_disparities=D, uniqueness_ratio=uniq, speckle_range=spr, speckle_window_size=spw,
                pre_filter_cap=cap, min_disparity=minD, outlier_min_k=k, outlier_std_threshold=alpha)


def _golden_inputs(O, g):
    H, W, D, bs, mode, minD, seed = [int(v) for v in g["meta"]]
    lb, rb = O.synth_pair(H, W, D, seed, color=True)
    assert zlib.crc32(lb.tobytes()) == int(g["crc"][0])
    return H, W, D, bs, mode, minD, lb, rb


@pytest.mark.parametrize("bs,mode", [(5, 0), (27, 0), (9, 1), (27, 1)])
def test_stages_against_oracle(cvr, oracle, monkeypatch, bs, mode):
    """Cost volume C, aggregated volume S, raw WTA/LR disparity and final disparity, each bit-exact."""
    monkeypatch.setenv("CVR_KEEP_S", "1")              # the last pass only writes the final S out when asked to
    H, W, D = 48, 260, 64
    l, r = oracle.synth_pair(H, W, D, seed=100 + bs + mode)
    p = oracle.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode)
    ref, st = oracle.sgbm(l, r, p, stages=True)
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D), mode=mode)
    got = rec.GenerateDisparityMap(l, r)
    C = rec.debug_fetch(0)[:, :, :D]
    S = rec.debug_fetch(1)[:, :, :D]
    lr = rec.debug_fetch(3)
    assert np.array_equal(C, st["C"]), "C mismatches: %d" % int((C != st["C"]).sum())
    assert np.array_equal(S, st["S"]), "S mismatches: %d" % int((S != st["S"]).sum())
    assert np.array_equal(lr, st["raw"]), "raw mismatches: %d" % int((lr != st["raw"]).sum())
    assert np.array_equal(got, ref), "final mismatches: %d" % int((got != ref).sum())
    rec.close()


@pytest.mark.parametrize("name", GOLDEN_SMALL + GOLDEN_LARGE)
def test_disparity_bit_exact_vs_cv2_golden(cvr, oracle, name):
    g = load_golden(name)
    H, W, D, bs, mode, minD, lb, rb = _golden_inputs(oracle, g)
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, minD), mode=mode)
    got = rec.GenerateDisparityMap(lb, rb)          # BGR in: cvtColor + StereoSGBM::compute
    assert np.array_equal(got, g["disp"]), "%d pixels differ from cv2" % int((got != g["disp"]).sum())
    rec.close()
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, minD, spw=0), mode=mode)
    got = rec.GenerateDisparityMap(lb, rb)
    assert np.array_equal(got, g["disp_nospeckle"])
    rec.close()


@pytest.mark.parametrize("name", GOLDEN_BIG)
def test_full_size_bit_exact_vs_cv2_golden(cvr, oracle, name):
    """BASELINE.json configs[2] / configs[3] at the reference's block size 27 (and 5): cv2's disparity, pruned disparity and
    prune mask by CRC32 plus every 16th disparity row.  hd_* is the shape that runs the vertical path kernel as a
    non-portable 16-CTA cluster with 512-thread CTAs (8 disparities per lane)."""
    g = load_golden(name)
    H, W, D, bs, mode, minD, lb, rb = _golden_inputs(oracle, g)
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, minD), mode=mode)
    disp = rec.GenerateDisparityMap(lb, rb)
    bad = int((disp[::16] != g["rows16"]).sum())
    assert bad == 0, "%d pixels of the sampled rows differ from cv2" % bad
    assert zlib.crc32(disp.tobytes()) == int(g["disp_crc"][0])
    assert int((disp >= 0).sum()) == int(g["valid_count"][0])
    pruned, dispf, mask = rec.SetupFrame(disp)
    assert zlib.crc32(pruned.tobytes()) == int(g["pruned_crc"][0]) and zlib.crc32(mask.tobytes()) == int(g["mask_crc"][0])
    assert np.array_equal(dispf, pruned.astype(np.float32) / np.float32(16))
    rec.close()


@pytest.mark.parametrize("name", GOLDEN_SMALL + GOLDEN_LARGE)
def test_prune_vs_cv2_golden(cvr, oracle, name):
    """TrackingFrame::SetupFrame on every fixture's cv2 disparity: pruned map and mask equal cv2's normalize / meanStdDev /
    threshold / convertTo chain."""
    g = load_golden(name)
    H, W, D, bs, mode, minD, seed = [int(v) for v in g["meta"]]
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, minD), mode=mode)
    pruned, dispf, mask = rec.SetupFrame(g["disp"])
    assert np.array_equal(pruned, g["pruned"]) and np.array_equal(mask, g["mask"])
    assert np.array_equal(dispf, g["pruned"].astype(np.float32) / np.float32(16))
    rec.close()


@pytest.mark.parametrize("mode", [0, 1])
def test_saturation_stick_and_drift_vs_cv2_golden(cvr, oracle, mode):
    g = load_golden("satdrift_d32_bs27")
    l, r = oracle.satdrift_pair()
    H, W = l.shape
    rec = cvr.Reconstruct3D(W, H, config=_cfg(27, 32), mode=mode)
    assert np.array_equal(rec.GenerateDisparityMap(l, r), g["disp_mode%d" % mode])
    rec.close()
    rec = cvr.Reconstruct3D(W, H, config=_cfg(27, 32, uniq=0, spw=0), mode=mode, P1=8, P2=32)
    assert np.array_equal(rec.GenerateDisparityMap(l, r), g["disp_smallP_mode%d" % mode])
    rec.close()


@pytest.mark.parametrize("kw", [dict(cap=63), dict(cap=31, uniq=0), dict(uniq=15, minD=16), dict(minD=-16, uniq=5), dict(D=16), dict(D=48, bs=1),
                                dict(D=96, bs=3), dict(D=160, bs=7), dict(D=192, bs=11), dict(D=256, bs=5), dict(D=112, bs=9), dict(D=32, bs=37), dict(D=80, bs=5, uniq=0)])
def test_parameter_sweep_against_oracle(cvr, oracle, kw):
    D = kw.get("D", 64)
    bs = kw.get("bs", 7)
    minD = kw.get("minD", 0)
    H, W = 40, D + 150
    for mode in (0, 1):
        l, r = oracle.synth_pair(H, W, D, seed=7 + D + bs)
        p = oracle.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode, min_disparity=minD,
                                         pre_filter_cap=kw.get("cap", 10), uniqueness_ratio=kw.get("uniq", 10))
        ref = oracle.sgbm(l, r, p)
        rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, minD, uniq=kw.get("uniq", 10), cap=kw.get("cap", 10)), mode=mode)
        got = rec.GenerateDisparityMap(l, r)
        assert np.array_equal(got, ref), "mode %d: %d pixels differ" % (mode, int((got != ref).sum()))
        rec.close()


@pytest.mark.parametrize("D,bs,mode", [(64, 5, 0), (128, 27, 0), (48, 9, 1), (256, 3, 1)])
def test_generic_int32_path_kernels(cvr, oracle, monkeypatch, D, bs, mode):
    """The int32 path kernels (used when P1 wraps negative or the shared-memory state does not fit) stay bit-exact;
    CVR_PATH_KERNELS=generic forces them."""
    monkeypatch.setenv("CVR_PATH_KERNELS", "generic")
    H, W = 44, D + 170
    l, r = oracle.synth_pair(H, W, D, seed=D + bs)
    ref = oracle.sgbm(l, r, oracle.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode))
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D), mode=mode)
    assert np.array_equal(rec.GenerateDisparityMap(l, r), ref)
    rec.close()
    # and on the adversarial saturation pair (negative C and negative path costs)
    g = load_golden("satdrift_d32_bs27")
    l, r = oracle.satdrift_pair()
    rec = cvr.Reconstruct3D(l.shape[1], l.shape[0], config=_cfg(27, 32), mode=mode)
    assert np.array_equal(rec.GenerateDisparityMap(l, r), g["disp_mode%d" % mode])
    rec.close()


@pytest.mark.parametrize("mode", [0, 1])
def test_no_neighbour_shortcut_on_and_off(cvr, oracle, monkeypatch, mode):
    """block_size 27 (P2 wraps to 4448 <= P1 = 17496) takes the no-neighbour form of the path update; with CVR_NO_NN=1
    the same parameters run through the ordinary packed update.  Both must equal cv2 / the oracle."""
    g = load_golden("s64x300_d64_bs27")
    H, W, D, bs, _, minD, lb, rb = _golden_inputs(oracle, g)
    gl, gr = oracle.gray_bgr(lb), oracle.gray_bgr(rb)
    ref = g["disp"] if mode == 0 else oracle.sgbm(gl, gr, oracle.reference_sgbm_params(block_size=bs, num_disparities=D, mode=1))
    for no_nn in ("0", "1"):
        monkeypatch.setenv("CVR_NO_NN", no_nn)
        rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, minD), mode=mode)
        assert np.array_equal(rec.GenerateDisparityMap(gl, gr), ref), "CVR_NO_NN=%s" % no_nn
        rec.close()
    # P1 >= P2s >= 0 with a P2 that does NOT wrap (P1 = 300, P2 = 301 after OpenCV's max(P2, P1+1)): also the shortcut regime
    l, r = oracle.synth_pair(40, 220, 32, seed=9)
    p = oracle.reference_sgbm_params(block_size=5, num_disparities=32, mode=mode)
    p["P1"], p["P2"] = 300, 10
    want = oracle.sgbm(l, r, p)
    for no_nn in ("0", "1"):
        monkeypatch.setenv("CVR_NO_NN", no_nn)
        rec = cvr.Reconstruct3D(220, 40, config=_cfg(5, 32), mode=mode, P1=300, P2=10)
        assert np.array_equal(rec.GenerateDisparityMap(l, r), want)
        rec.close()


@pytest.mark.parametrize("W1", [1, 3, 7, 8, 9, 17, 33, 65, 127])
def test_narrow_images_pick_every_cluster_geometry(cvr, oracle, W1):
    """Very narrow computed ranges walk through all launch geometries of the vertical path kernel (cluster size 1..8,
    256..1024 threads, down to the int32 fallback when even 8 warps do not get a column each)."""
    D, H = 32, 36
    W = D + W1
    for bs, mode in ((5, 0), (27, 1)):
        l, r = oracle.synth_pair(H, W, D, seed=W1 + bs)
        want = oracle.sgbm(l, r, oracle.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode))
        rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D), mode=mode)
        assert np.array_equal(rec.GenerateDisparityMap(l, r), want), (W1, bs, mode)
        rec.close()


def test_degenerate_inputs(cvr, oracle):
    # image narrower than the disparity range: everything invalid (A.0, W1 <= 0)
    rec = cvr.Reconstruct3D(60, 20, config=_cfg(5, 64))
    z = np.zeros((20, 60), np.uint8)
    assert (rec.GenerateDisparityMap(z, z) == -16).all()
    rec.close()
    # constant images
    rec = cvr.Reconstruct3D(200, 30, config=_cfg(5, 32))
    a = np.full((30, 200), 77, np.uint8)
    p = oracle.reference_sgbm_params(block_size=5, num_disparities=32)
    assert np.array_equal(rec.GenerateDisparityMap(a, a), oracle.sgbm(a, a, p))
    # padded row step (cv::Mat step > cols) through the raw ABI
    import ctypes as C
    lp = np.zeros((30, 256), np.uint8)
    rp = np.zeros((30, 256), np.uint8)
    l, r = oracle.synth_pair(30, 200, 32, 3)
    lp[:, :200], rp[:, :200] = l, r
    out = np.empty((30, 200), np.int16)
    rc = rec.lib.cvr_disparity(rec.ctx, lp.ctypes.data, 256, rp.ctypes.data, 256, 1, out.ctypes.data, 400)
    assert rc == 0 and np.array_equal(out, oracle.sgbm(l, r, p))
    # wrong channel count is an error, like cv::cvtColor throwing
    assert rec.lib.cvr_disparity(rec.ctx, lp.ctypes.data, 256, rp.ctypes.data, 256, 2, out.ctypes.data, 400) == -1
    rec.close()


def test_prune_reproject_triangulate(cvr, oracle):
    g = load_golden("s100x420_d96_bs3")
    H, W, D, bs, mode, minD, lb, rb = _golden_inputs(oracle, g)
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, minD), mode=mode)
    pruned, dispf, mask = rec.SetupFrame(g["disp"])
    assert np.array_equal(pruned, g["pruned"]) and np.array_equal(mask, g["mask"])          # vs cv2
    assert np.array_equal(dispf, g["pruned"].astype(np.float32) / np.float32(16))
    cloud = rec.GeneratePointCloud(pruned, lb)
    ref = oracle.generate_point_cloud(pruned, lb, rec.Q)
    assert len(cloud) == len(ref) == int(g["xyz_valid_count"][0])
    for f in ("x", "y", "z"):
        assert np.array_equal(cloud[f].view(np.uint32), ref[f].view(np.uint32)), f      # bitwise (bar: 1e-5 relative)
    assert np.array_equal(cloud["r"], ref["r"]) and np.array_equal(cloud["b"], ref["b"])
    tri = rec.Triangulate3D(dispf, lb, mask)
    p = rec.params
    tref = oracle.triangulate(dispf, lb, mask, p.tri_f, p.tri_b, p.tri_cx, p.tri_cy)
    assert len(tri) == len(tref) > 1000
    for f in ("x", "y", "z"):
        assert np.allclose(tri[f], tref[f], rtol=1e-5, atol=0)
        assert np.array_equal(tri[f].view(np.uint32), tref[f].view(np.uint32)), f
    # no mask: every positive disparity
    tri2 = rec.Triangulate3D(dispf, lb, None)
    assert len(tri2) == int((dispf > 0).sum())
    rec.close()


def _cloud_for_filters(cvr, oracle, name="s100x420_d96_bs3"):
    g = load_golden(name)
    H, W, D, bs, mode, minD, lb, rb = _golden_inputs(oracle, g)
    pruned, dispf, mask, _, _ = oracle.prune(g["disp"])
    pts = oracle.triangulate(dispf, lb, mask, np.float32(958.35805), np.float32(0.532671), np.float32(W / 2), np.float32(H / 2))
    return H, W, pts


def test_voxel_grid_vs_oracle(cvr, oracle):
    H, W, pts = _cloud_for_filters(cvr, oracle)
    for leaf in (0.2, 0.05):
        rec = cvr.Reconstruct3D(W, H, config=_cfg(3, 96), voxel_leaf=leaf)
        got, pt = rec.VoxelGrid(pts)
        ref, counts, rpt = oracle.voxel_grid(pts, leaf)
        assert pt == rpt and len(got) == len(ref)
        for f in ("x", "y", "z"):
            assert np.abs(got[f] - ref[f]).max() <= 1e-5 * leaf
        for f in ("r", "g", "b"):
            assert np.abs(got[f].astype(int) - ref[f].astype(int)).max() <= 1
        rec.close()
    # int32 index overflow -> PCL passes the cloud through unchanged
    rec = cvr.Reconstruct3D(W, H, config=_cfg(3, 96), voxel_leaf=1e-6)
    got, pt = rec.VoxelGrid(pts)
    assert pt and np.array_equal(got, pts)
    assert len(rec.VoxelGrid(pts[:0])[0]) == 0
    rec.close()


def test_sor_vs_oracle(cvr, oracle):
    H, W, pts = _cloud_for_filters(cvr, oracle)
    vox, _, _ = oracle.voxel_grid(pts, 0.05)
    assert len(vox) > 500
    for (k, alpha) in ((30, 0.5), (8, 1.0), (50, 0.3)):
        rec = cvr.Reconstruct3D(W, H, config=_cfg(3, 96, k=k, alpha=alpha))
        got = rec.RemoveOutliers(vox)
        ref, dist, thr, keep = oracle.sor(vox, k, alpha)
        near = np.abs(dist - thr) <= 1e-5 * thr
        sure = vox[(keep == 1) & ~near]
        # every surely-kept point is present, in input order, and nothing surely-removed appears
        gk = set(map(bytes, got))
        assert all(bytes(p) in gk for p in sure)
        assert len(got) <= int(keep.sum()) + int(near.sum()) and len(got) >= len(sure)
        if not near.any():
            assert np.array_equal(got, ref)
        rec.close()
    # fewer points than k+1: nothing can be judged, everything is kept
    rec = cvr.Reconstruct3D(W, H, config=_cfg(3, 96, k=30))
    assert np.array_equal(rec.RemoveOutliers(vox[:10]), vox[:10])
    # duplicates are legal
    dup = np.concatenate([vox[:200], vox[:200]])
    assert np.array_equal(rec.RemoveOutliers(dup), oracle.sor(dup, 30, 0.5)[0])
    rec.close()


def test_batch_equals_single_frames(cvr, oracle):
    H, W, D, bs = 60, 330, 64, 9
    B = 5
    Ls, Rs = zip(*[oracle.synth_pair(H, W, D, seed=40 + i, color=True) for i in range(B)])
    Ls, Rs = np.stack(Ls), np.stack(Rs)
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, k=8), max_batch=B, voxel_leaf=0.05)
    disp, clouds = rec.ProcessFrames(Ls, Rs, want_disp=True)
    p = oracle.reference_sgbm_params(block_size=bs, num_disparities=D)
    prm = rec.params
    for i in range(B):
        ref = oracle.sgbm(oracle.gray_bgr(Ls[i]), oracle.gray_bgr(Rs[i]), p)
        assert np.array_equal(disp[i], ref), "frame %d" % i
        pruned, dispf, mask, _, _ = oracle.prune(ref)
        pts = oracle.triangulate(dispf, Ls[i], mask, prm.tri_f, prm.tri_b, prm.tri_cx, prm.tri_cy)
        vox, _, _ = oracle.voxel_grid(pts, 0.05)
        kept, dist, thr, keep = oracle.sor(vox, 8, 0.5)
        near = int((np.abs(dist - thr) <= 1e-5 * thr).sum())
        assert abs(len(clouds[i]) - len(kept)) <= near
        if near == 0:
            assert np.array_equal(clouds[i]["x"], kept["x"]) and np.array_equal(clouds[i]["r"], kept["r"])
    rec.close()


@pytest.mark.parametrize("bs", [3, 9, 19])
def test_mode_hh_second_pass_order_with_mixed_batches(cvr, oracle, monkeypatch, bs):
    """MODE_HH: frames whose regime flags stay clear take the cheaper order of the second pass (upward paths, then <- with the
    winner-take-all), frames with a flag (negative costs: small or wrapping P2, saturating input) are redone in OpenCV's order.
    A batch that mixes textured, saturating, noise and constant frames must equal the oracle frame by frame in both settings,
    and with the fallback kernels."""
    H, W, D, B = 48, 300, 64, 6
    rng = np.random.default_rng(bs)
    Ls, Rs = [], []
    for i in range(B):
        l, r = oracle.synth_pair(H, W, D, seed=70 + i, color=True)
        if i == 1:                                               # saturating: black / white stripes
            l = np.where((np.arange(W) // 7 % 2)[None, :, None] > 0, 255, 0).astype(np.uint8) * np.ones((H, 1, 3), np.uint8); r = np.roll(l, -9, axis=1)
        elif i == 3:
            l = rng.integers(0, 256, (H, W, 3), dtype=np.uint8); r = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        elif i == 4:
            l = np.full((H, W, 3), 90, np.uint8); r = l.copy()
        Ls.append(np.ascontiguousarray(l)); Rs.append(np.ascontiguousarray(r))
    Ls, Rs = np.stack(Ls), np.stack(Rs)
    p = oracle.reference_sgbm_params(block_size=bs, num_disparities=D, mode=1, pre_filter_cap=63)
    want = [oracle.sgbm(oracle.gray_bgr(Ls[i]), oracle.gray_bgr(Rs[i]), p) for i in range(B)]
    for env in ({}, {"CVR_HH_ORDER": "exact"}, {"CVR_PATH_KERNELS": "v3"}):
        for k in ("CVR_HH_ORDER", "CVR_PATH_KERNELS"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, cap=63), mode=1, max_batch=B)
        disp, _ = rec.ProcessFrames(Ls, Rs, want_disp=True, stages=1)
        for i in range(B):
            assert np.array_equal(disp[i], want[i]), (env, i)
        rec.close()


def test_pipelined_submit_wait_equals_synchronous(cvr, oracle):
    """cvr_submit_batch / cvr_wait_batch (two batches in flight) return exactly what cvr_process_batch returns."""
    H, W, D, bs, B, NB = 60, 330, 64, 9, 3, 4
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, k=8), max_batch=B, voxel_leaf=0.05)
    batches = []
    for nb in range(NB):
        Ls, Rs = zip(*[oracle.synth_pair(H, W, D, seed=70 + nb * B + i, color=True) for i in range(B)])
        batches.append((np.stack(Ls), np.stack(Rs)))
    want = [rec.ProcessFrames(L, R, want_disp=True) for (L, R) in batches]
    cap = W * H
    ios, outs = [], []
    for (L, R) in batches:
        disp = np.empty((B, H, W), np.int16)
        clouds = np.empty((B, cap), cvr.POINT_DTYPE)
        counts = np.zeros(B, np.int32)
        io = cvr.Batch()
        io.mem, io.channels = cvr.MEM_HOST, 3
        io.left, io.left_step, io.left_frame_stride = L.ctypes.data, L.strides[1], L.strides[0]
        io.right, io.right_step, io.right_frame_stride = R.ctypes.data, R.strides[1], R.strides[0]
        io.disp, io.disp_step, io.disp_frame_stride = disp.ctypes.data, disp.strides[1], disp.strides[0]
        io.cloud, io.cloud_capacity, io.cloud_count = clouds.ctypes.data, cap, counts.ctypes.data
        io.stages = 15
        ios.append(io)
        outs.append((disp, clouds, counts))
    rec.submit_batch_raw(ios[0], B)
    for nb in range(1, NB):
        rec.submit_batch_raw(ios[nb], B)
        rec.wait_batch()
    with pytest.raises(cvr.CvrError):          # only one batch is left in flight: a second wait after the last must fail
        rec.wait_batch()
        rec.wait_batch()
    for nb in range(NB):
        disp, clouds, counts = outs[nb]
        wd, wc = want[nb]
        assert np.array_equal(disp, wd), "batch %d" % nb
        for i in range(B):
            assert counts[i] == len(wc[i]) and np.array_equal(clouds[i, :counts[i]], wc[i])
    # three outstanding submits are refused
    rec.submit_batch_raw(ios[0], B)
    rec.submit_batch_raw(ios[1], B)
    with pytest.raises(cvr.CvrError):
        rec.submit_batch_raw(ios[2], B)
    rec.wait_batch()
    rec.wait_batch()
    rec.close()


@pytest.mark.parametrize("rect", [False, True])
def test_pipelined_device_resident_batches(cvr, oracle, rect):
    """cvr_submit_batch with DEVICE buffers: inputs are read in place, outputs land in the caller's device buffers; the
    matcher of batch i+1 overlaps the filters of batch i (with rectification: shares the rectified-image buffer)."""
    import torch
    H, W, D, bs, B, NB = 60, 330, 64, 9, 3, 5
    cfg = _cfg(bs, D, k=8)
    cfg["requires_rectification"] = rect
    rec = cvr.Reconstruct3D(W, H, config=cfg, max_batch=B, voxel_leaf=0.05)
    if rect:
        K = np.array([[300.0, 0, 165.0], [0, 300.0, 30.0], [0, 0, 1]])
        P = np.array([[295.0, 0, 163.0, 0], [0, 295.0, 31.0, 0], [0, 0, 1, 0]])
        rec.SetRectification(K, [-0.05, 0.01, 0.0003, -0.0002, 0.0], np.eye(3), P, K, [-0.04, 0.012, -0.0001, 0.0002, 0.0], np.eye(3), P)
    batches = []
    for nb in range(NB):
        Ls, Rs = zip(*[oracle.synth_pair(H, W, D, seed=170 + nb * B + i, color=True) for i in range(B)])
        batches.append((np.stack(Ls), np.stack(Rs)))
    want = [rec.ProcessFrames(L, R, want_disp=True) for (L, R) in batches]
    cap = W * H
    dev = torch.device("cuda", 0)
    ios, outs, keep = [], [], []
    for (L, R) in batches:
        dL, dR = torch.from_numpy(L).to(dev), torch.from_numpy(R).to(dev)
        disp = torch.empty((B, H, W), dtype=torch.int16, device=dev)
        clouds = torch.zeros((B, cap, 16), dtype=torch.uint8, device=dev)
        counts = torch.zeros(B, dtype=torch.int32, device=dev)
        io = cvr.Batch()
        io.mem, io.channels = cvr.MEM_DEVICE, 3
        io.left, io.left_step, io.left_frame_stride = dL.data_ptr(), W * 3, W * H * 3
        io.right, io.right_step, io.right_frame_stride = dR.data_ptr(), W * 3, W * H * 3
        io.disp, io.disp_step, io.disp_frame_stride = disp.data_ptr(), W * 2, W * H * 2
        io.cloud, io.cloud_capacity, io.cloud_count = clouds.data_ptr(), cap, counts.data_ptr()
        io.stages = 15
        ios.append(io); outs.append((disp, clouds, counts)); keep.append((dL, dR))
    torch.cuda.synchronize()
    rec.submit_batch_raw(ios[0], B)
    for nb in range(1, NB):
        rec.submit_batch_raw(ios[nb], B)
        rec.wait_batch()
    rec.wait_batch()
    for nb in range(NB):
        disp, clouds, counts = (t.cpu().numpy() for t in outs[nb])
        wd, wc = want[nb]
        assert np.array_equal(disp, wd), "batch %d" % nb
        for i in range(B):
            got = clouds[i, :counts[i]].reshape(-1).view(cvr.POINT_DTYPE)
            assert counts[i] == len(wc[i]) and np.array_equal(got, wc[i]), "batch %d frame %d" % (nb, i)
    rec.close()


@pytest.mark.parametrize("extra", [5, 6, 7, 8])
def test_unaligned_device_images_all_width_residues(cvr, oracle, extra):
    """Device-resident BGR frames with a padded row step and a base pointer that is not 4-byte aligned, image widths of
    every residue mod 4 (the gray kernel reads aligned words and funnel-shifts; the median takes pixel pairs only for
    even widths)."""
    import torch
    D, bs, H, B = 32, 5, 11, 3
    W = D + 60 + extra
    pairs = [oracle.synth_pair(H, W, D, seed=300 + extra * 10 + i, color=True) for i in range(B)]
    step, off = W * 3 + 5, 1
    fs = step * H + 3
    dev = torch.device("cuda", 0)
    bufL = torch.zeros(off + fs * B + 64, dtype=torch.uint8, device=dev)
    bufR = torch.zeros_like(bufL)
    for i, (l, r) in enumerate(pairs):
        for (buf, img) in ((bufL, l), (bufR, r)):
            view = buf[off + i * fs: off + i * fs + step * H].view(H, step)
            view[:, :W * 3] = torch.from_numpy(img.reshape(H, W * 3)).to(dev)
    disp = torch.empty((B, H, W), dtype=torch.int16, device=dev)
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D), max_batch=B)
    io = cvr.Batch()
    io.mem, io.channels = cvr.MEM_DEVICE, 3
    io.left, io.left_step, io.left_frame_stride = bufL.data_ptr() + off, step, fs
    io.right, io.right_step, io.right_frame_stride = bufR.data_ptr() + off, step, fs
    io.disp, io.disp_step, io.disp_frame_stride = disp.data_ptr(), W * 2, W * H * 2
    io.stages = 1
    torch.cuda.synchronize()
    rec.process_batch_raw(io, B)
    got = disp.cpu().numpy()
    p = oracle.reference_sgbm_params(block_size=bs, num_disparities=D)
    for i, (l, r) in enumerate(pairs):
        assert np.array_equal(got[i], oracle.sgbm(oracle.gray_bgr(l), oracle.gray_bgr(r), p)), "frame %d" % i
    rec.close()


def test_suggest_batch_fills_cluster_waves(cvr):
    """cvr_suggest_batch: stays within a quarter of the target and below max_batch, is idempotent, and leaves the target
    alone where the clustered kernel is not used (StereoBM)."""
    rec = cvr.Reconstruct3D(1242, 375, config=cvr.parse_config(None), max_batch=40)
    b = rec.suggest_batch(30)
    assert 23 <= b <= 37 and rec.suggest_batch(b) == b
    assert rec.suggest_batch(1000) <= 40 and rec.suggest_batch(1) == 1
    rec.close()
    cfg = cvr.parse_config(None)
    cfg["block_matcher"] = cvr.MATCHER_BM
    rec = cvr.Reconstruct3D(640, 120, config=cfg, max_batch=16)
    assert rec.suggest_batch(9) == 9
    rec.close()


def _bm_cfg(p):
    return dict(block_matcher=1, window_size=p["blockSize"], bm_num_disparities=p["numDisparities"], bm_min_disparity=p["minDisparity"],
                bm_pre_filter_cap=p["preFilterCap"], bm_texture_threshold=p["textureThreshold"], bm_uniqueness_ratio=p["uniquenessRatio"],
                bm_speckle_window_size=p["speckleWindowSize"], bm_speckle_range=p["speckleRange"],
                block_size=5, num_disparities=16, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10, min_disparity=0)


@pytest.mark.parametrize("name", ["bm_60x300_d64_w9", "bm_61x300_d64_w21", "bm_64x260_d32_w11_minneg", "bm_64x260_d48_w15_nouniq",
                                  "bm_64x260_d48_w15_speckle", "bm_kitti_d128_w27"])
def test_stereobm_bit_exact_vs_cv2_golden(cvr, oracle, name):
    """The reference's default block matcher (config_default.json "block_matcher": "stereo_bm"), SURVEY 8(f) n2."""
    from test_oracle_golden import bm_golden_inputs
    g = load_golden(name)
    lb, rb, p = bm_golden_inputs(oracle, g)
    H, W = lb.shape[:2]
    rec = cvr.Reconstruct3D(W, H, config=_bm_cfg(p))
    got = rec.GenerateDisparityMap(lb, rb)
    assert np.array_equal(got, g["disp"]), "%d pixels differ from cv2.StereoBM" % int((got != g["disp"]).sum())
    rec.close()


@pytest.mark.parametrize("D,ws", [(16, 5), (32, 7), (80, 13), (160, 9), (256, 17), (96, 31)])
def test_stereobm_sweep_against_oracle(cvr, oracle, D, ws):
    H, W = 50 + ws, D + 140
    l, r = oracle.synth_pair(H, W, D, seed=D + ws)
    p = oracle.reference_bm_params(ws, D)
    want = oracle.stereobm(l, r, p)
    rec = cvr.Reconstruct3D(W, H, config=_bm_cfg(p), max_batch=2)
    assert np.array_equal(rec.GenerateDisparityMap(l, r), want)
    # through the batched chain as well (disparity -> prune -> cloud ...)
    disp, clouds = rec.ProcessFrames(np.stack([l, l]), np.stack([r, r]), want_disp=True, stages=3)
    assert np.array_equal(disp[0], want) and np.array_equal(disp[1], want)
    rec.close()


def test_stereobm_rejects_what_opencv_rejects(cvr):
    for bad in (dict(window_size=4), dict(window_size=3), dict(bm_num_disparities=24), dict(bm_pre_filter_cap=64), dict(bm_min_disparity=8)):
        cfg = _bm_cfg(dict(blockSize=9, numDisparities=32, minDisparity=0, preFilterCap=31, textureThreshold=10, uniquenessRatio=15,
                           speckleWindowSize=0, speckleRange=0))
        cfg.update(bad)
        with pytest.raises(cvr.CvrError):
            cvr.Reconstruct3D(200, 60, config=cfg)


def test_radius_outlier_removal_and_transform_vs_oracle(cvr, oracle):
    """pcl::RadiusOutlierRemoval and pcl::transformPointCloud as MappingSystem.cpp:303-318 calls them."""
    H, W, pts = _cloud_for_filters(cvr, oracle)
    vox, _, _ = oracle.voxel_grid(pts, 0.05)
    rec = cvr.Reconstruct3D(W, H, config=_cfg(3, 96))
    for radius, mn in ((0.2, 5), (0.5, 40), (1.0, 300), (0.05, 0)):
        got = rec.RadiusOutlierRemoval(vox, radius, mn)
        ref, keep = oracle.radius_outlier_removal(vox, radius, mn)
        assert np.array_equal(got, ref), (radius, mn, len(got), len(ref))
    assert 0 < len(oracle.radius_outlier_removal(vox, 0.2, 5)[0]) < len(vox)            # the comparison above is not vacuous
    th = 0.3
    T = np.array([[np.cos(th), 0, np.sin(th), 1.5], [0, 1, 0, -0.25], [-np.sin(th), 0, np.cos(th), 7.0], [0, 0, 0, 1]], np.float32)
    got = rec.TransformPointCloud(vox, T)
    ref = oracle.transform_cloud(vox, T)
    for f in ("x", "y", "z"):
        assert np.array_equal(got[f].view(np.uint32), ref[f].view(np.uint32))
    assert np.array_equal(got["r"], vox["r"])
    rec.close()
