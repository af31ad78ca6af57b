"""GPU: stereo rectification of the inputs (Reconstruct3D::RectifyImages, Reconstruct3D.cpp:330-361)
through the C ABI against the cv2 fixtures (tests/golden/rect_*.npz) and the CPU oracle.  Bar: bit-exact maps
(CV_16SC2 + CV_16UC1) and bit-exact remapped images."""
import zlib

import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu

RECT_GOLDEN = ["rect_small_k5", "rect_small_k8_border", "rect_small_nodist_shift", "rect_small_k12", "rect_kitti_k5"]


def rect_image(H, W, ch, seed):
    img = np.random.default_rng(seed).integers(0, 256, (H, W, ch), dtype=np.uint8)
    return img[:, :, 0].copy() if ch == 1 else img


def _cfg(bs, D, rect=False):
    return dict(block_size=bs, num_disparities=D, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10,
                min_disparity=0, outlier_min_k=8, outlier_std_threshold=0.5, requires_rectification=rect)


@pytest.mark.parametrize("name", RECT_GOLDEN)
def test_rectification_bit_exact_vs_cv2_golden(cvr, oracle, name):
    g = load_golden(name)
    H, W, ch, seed = (int(v) for v in g["meta"])
    img = rect_image(H, W, ch, seed)
    assert zlib.crc32(img.tobytes()) == int(g["crc"][0]), "input generator drifted"
    rec = cvr.Reconstruct3D(W, H, config=_cfg(5, 16))
    # camera 1 takes the fixture's calibration, camera 2 the same with the image mirrored in the test below
    rec.SetRectification(g["K"], g["D"], g["R"], g["P"], g["K"], g["D"], g["R"], g["P"])
    for cam in (0, 1):
        m1, m2 = rec.rectification_maps(cam)
        assert zlib.crc32(m1.tobytes()) == int(g["crc"][1]), "map1 of camera %d differs from cv2" % cam
        assert zlib.crc32(m2.tobytes()) == int(g["crc"][2]), "map2 of camera %d differs from cv2" % cam
    other = np.ascontiguousarray(img[::-1])                 # a different right image through the same maps
    outL, outR = rec.RectifyImages(img, other)
    assert zlib.crc32(outL.tobytes()) == int(g["crc"][3]), "remapped image differs from cv2.remap"
    if "out" in g.files:
        assert np.array_equal(outL, g["out"])
    om1, om2 = oracle.init_undistort_rectify_map(g["K"], g["D"], g["R"], g["P"], W, H)
    assert np.array_equal(outR, oracle.remap_bilinear(other, om1, om2))
    rec.close()


def test_rectification_two_cameras_and_channel_counts(cvr, oracle):
    """Different calibrations per camera; 1-, 3- and 4-channel images; maps that leave the image on every side."""
    H, W = 70, 257
    K1 = np.array([[210.0, 0, 128.0], [0, 212.0, 36.0], [0, 0, 1]])
    K2 = np.array([[205.0, 0, 131.0], [0, 207.0, 33.0], [0, 0, 1]])
    D1 = [-0.31, 0.12, 0.001, -0.002, -0.03]
    D2 = [-0.05, 0.01, 0.0, 0.0]
    th = 0.12
    R1 = np.array([[np.cos(th), -np.sin(th), 0], [np.sin(th), np.cos(th), 0], [0, 0, 1]])       # in-plane rotation: corners leave the image
    R2 = np.eye(3)
    P1 = np.array([[150.0, 0, 128.0, 0], [0, 150.0, 35.0, 0], [0, 0, 1, 0]])                  # zoomed out: a border all around
    P2 = np.array([[260.0, 0, 100.0, -40.0], [0, 260.0, 50.0, 0], [0, 0, 1, 0]])
    rec = cvr.Reconstruct3D(W, H, config=_cfg(5, 16))
    rec.SetRectification(K1, D1, R1, P1, K2, D2, R2, P2)
    o = [oracle.init_undistort_rectify_map(K1, D1, R1, P1, W, H), oracle.init_undistort_rectify_map(K2, D2, R2, P2, W, H)]
    for cam in (0, 1):
        m1, m2 = rec.rectification_maps(cam)
        assert np.array_equal(m1, o[cam][0]) and np.array_equal(m2, o[cam][1])
    assert (o[0][0][..., 0] < -1).any() and (o[0][0][..., 0] >= W).any() and (o[0][0][..., 1] < -1).any() and (o[0][0][..., 1] >= H).any()
    for ch in (1, 3, 4):
        l, r = rect_image(H, W, ch, 50 + ch), rect_image(H, W, ch, 60 + ch)
        gl, gr = rec.RectifyImages(l, r)
        assert np.array_equal(gl, oracle.remap_bilinear(l, *o[0])), "left, %d channels" % ch
        assert np.array_equal(gr, oracle.remap_bilinear(r, *o[1])), "right, %d channels" % ch
    rec.close()


def test_requires_rectification_in_the_batch_path(cvr, oracle):
    """config "requires_rectification": true (ConfigParser.cpp:39, ReconstructionSystem.cpp:42-44): the batch is rectified
    on the device first; disparity and cloud must equal rectifying with the oracle and processing the result."""
    H, W, D, bs, B = 64, 320, 32, 7, 3
    K = np.array([[300.0, 0, 160.0], [0, 300.0, 32.0], [0, 0, 1]])
    Dl, Dr = [-0.08, 0.02, 0.0005, -0.0004, 0.0], [-0.06, 0.015, -0.0003, 0.0002, 0.0]
    R = np.eye(3)
    P = np.array([[290.0, 0, 158.0, 0], [0, 290.0, 33.0, 0], [0, 0, 1, 0]])
    pairs = [oracle.synth_pair(H, W, D, seed=70 + i, color=True) for i in range(B)]
    Ls, Rs = np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])
    rec = cvr.Reconstruct3D(W, H, config=_cfg(bs, D, rect=True), max_batch=B, voxel_leaf=0.05)
    with pytest.raises(cvr.CvrError):
        rec.ProcessFrames(Ls, Rs, want_disp=True)             # maps not set yet: refused, not silently skipped
    rec.SetRectification(K, Dl, R, P, K, Dr, R, P)
    disp, clouds = rec.ProcessFrames(Ls, Rs, want_disp=True)
    plain = cvr.Reconstruct3D(W, H, config=_cfg(bs, D), max_batch=B, voxel_leaf=0.05)
    mL, mR = oracle.init_undistort_rectify_map(K, Dl, R, P, W, H), oracle.init_undistort_rectify_map(K, Dr, R, P, W, H)
    rl = np.stack([oracle.remap_bilinear(Ls[i], *mL) for i in range(B)])
    rr = np.stack([oracle.remap_bilinear(Rs[i], *mR) for i in range(B)])
    disp2, clouds2 = plain.ProcessFrames(rl, rr, want_disp=True)
    p = oracle.reference_sgbm_params(block_size=bs, num_disparities=D)
    for i in range(B):
        assert np.array_equal(disp[i], oracle.sgbm(oracle.gray_bgr(rl[i]), oracle.gray_bgr(rr[i]), p)), "frame %d" % i
        assert np.array_equal(disp[i], disp2[i])
        assert len(clouds[i]) == len(clouds2[i]) and clouds[i].tobytes() == clouds2[i].tobytes()
    rec.close(); plain.close()


def test_rectification_rejects_what_it_does_not_implement(cvr):
    rec = cvr.Reconstruct3D(64, 48, config=_cfg(5, 16))
    K = np.eye(3); R = np.eye(3); P = np.eye(3)
    with pytest.raises(cvr.CvrError):
        rec.SetRectification(K, np.zeros(14), R, P, K, [], R, P)         # tilt terms
    with pytest.raises(cvr.CvrError):
        rec.RectifyImages(np.zeros((48, 64, 3), np.uint8), np.zeros((48, 64, 3), np.uint8))   # before SetRectification
    rec.close()
