"""Generates tests/golden/*.npz with the REAL OpenCV (python cv2) -- the library the reference
calls at Reconstruct3D.cpp:78-81,109 and TrackingFrame.cpp:32,51-54 -- on seeded synthetic pairs.

Run in the build container (cv2 4.13.0):  python tests/golden/make_golden.py
The fixtures pin the CPU oracle (tests/test_oracle_golden.py, CPU) and the CUDA path
(tests/test_gpu_parity.py, GPU).  Inputs are regenerated from oracle.synth_pair(seed) at test time;
each fixture stores a CRC of the inputs so generator drift is detected.
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cv2  # noqa: E402

from oracle import oracle as O  # noqa: E402

cv2.setNumThreads(1)

# name: (H, W, D, block_size, mode, minDisparity, seed, extras)
CASES = {
    "s64x300_d64_bs5": (64, 300, 64, 5, 0, 0, 11),
    "s64x300_d64_bs27": (64, 300, 64, 27, 0, 0, 12),
    "s120x500_d64_bs27_hh": (120, 500, 64, 27, 1, 0, 13),
    "s90x400_d48_bs9_minneg": (90, 400, 48, 9, 0, -16, 14),
    "s90x400_d32_bs7_hh_minpos": (90, 400, 32, 7, 1, 16, 15),
    "s100x420_d96_bs3": (100, 420, 96, 3, 0, 0, 16),
    "kitti_d128_bs27": (375, 1242, 128, 27, 0, 0, 0),
    "kitti_d128_bs9": (375, 1242, 128, 9, 0, 0, 1),
    "spi_d192_bs5_hh": (720, 1280, 192, 5, 1, 0, 2),
}

KITTI_F_REPROJECT = np.float32((959.791 + 903.7596) / 2.0)   # (K_L(0,0)+K_R(0,0))/2, Reconstruct3D.cpp:33


def reference_Q():
    """Reconstruct3D ctor (Reconstruct3D.cpp:31-35): Q = I4, Q(2,2) = 0.005*f (double product stored f32)."""
    Q = np.eye(4, dtype=np.float32)
    Q[2, 2] = np.float32(0.005 * float(KITTI_F_REPROJECT))
    return Q


def cv_sgbm(l, r, p, speckle=True):
    m = cv2.StereoSGBM_create(
        p["minDisparity"], p["numDisparities"], p["blockSize"], P1=p["P1"], P2=p["P2"],
        disp12MaxDiff=p["disp12MaxDiff"], preFilterCap=p["preFilterCap"], uniquenessRatio=p["uniquenessRatio"],
        speckleWindowSize=p["speckleWindowSize"] if speckle else 0, speckleRange=p["speckleRange"],
        mode=cv2.STEREO_SGBM_MODE_HH if p["mode"] == 1 else cv2.STEREO_SGBM_MODE_SGBM)
    return m.compute(l, r)


def main():
    for name, (H, W, D, bs, mode, minD, seed) in CASES.items():
        lb, rb = O.synth_pair(H, W, D, seed, color=True)
        gl = cv2.cvtColor(lb, cv2.COLOR_BGR2GRAY)
        gr = cv2.cvtColor(rb, cv2.COLOR_BGR2GRAY)
        p = O.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode, min_disparity=minD)
        disp = cv_sgbm(gl, gr, p, True)
        disp_nospk = cv_sgbm(gl, gr, p, False)
        # TrackingFrame::SetupFrame with cv2
        dn = cv2.normalize(disp, None, 0, 255, cv2.NORM_MINMAX, cv2.CV_8U)
        _, std = cv2.meanStdDev(dn)
        _, mask = cv2.threshold(dn, std[0][0] * 1.3, 255, cv2.THRESH_BINARY)
        pruned = disp.copy()
        pruned[mask == 0] = 0
        dispf = np.empty(pruned.shape, np.float32)
        cv2.Mat  # (cv2.Mat exists in >= 4.5; convertTo is spelled via multiply below)
        dispf = cv2.multiply(pruned.astype(np.float32), 1.0 / 16.0)
        mask2 = mask.copy()
        mask2[dispf <= 0] = 0
        xyz = cv2.reprojectImageTo3D(pruned, reference_Q(), handleMissingValues=True)
        out = dict(
            meta=np.array([H, W, D, bs, mode, minD, seed], np.int64),
            crc=np.array([zlib.crc32(lb.tobytes()), zlib.crc32(rb.tobytes())], np.uint64),
            gray_crc=np.array([zlib.crc32(gl.tobytes()), zlib.crc32(gr.tobytes())], np.uint64),
            disp=disp, disp_nospeckle=disp_nospk, pruned=pruned, mask=mask2, sigma=np.float64(std[0][0]),
            xyz_crc=np.array([zlib.crc32(xyz.tobytes())], np.uint64),
            xyz_valid_count=np.array([int((xyz[..., 2] != 10000.0).sum())], np.int64),
            cv_version=np.array(cv2.__version__),
        )
        if H * W < 100000:
            out["gray_left"] = gl
            out["xyz"] = xyz
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "valid %.1f%%" % (100 * (disp >= minD * 16).mean()), os.path.getsize(os.path.join(HERE, name + ".npz")) // 1024, "KiB")


def special():
    """Saturate-and-drift case: cost volume sticks at 32767, then leaves saturation (pins the A.3 recurrences)."""
    l, r = O.satdrift_pair()
    H, W = l.shape
    out = {}
    for mode in (0, 1):
        p = O.reference_sgbm_params(block_size=27, num_disparities=32, mode=mode)
        out["disp_mode%d" % mode] = cv_sgbm(l, r, p, True)
        p5 = O.reference_sgbm_params(block_size=27, num_disparities=32, mode=mode, uniqueness_ratio=0, speckle_window_size=0)
        p5["P1"], p5["P2"] = 8, 32          # small penalties: everything else unsaturated, only C saturates
        out["disp_smallP_mode%d" % mode] = cv_sgbm(l, r, p5, False)
    out["crc"] = np.array([zlib.crc32(l.tobytes()), zlib.crc32(r.tobytes())], np.uint64)
    np.savez_compressed(os.path.join(HERE, "satdrift_d32_bs27.npz"), **out)
    print("satdrift", {k: int((v >= 0).sum()) for k, v in out.items() if k.startswith("disp")})


BM_CASES = {   # name: (H, W, ndisp, window, seed, extra StereoBM setters)
    "bm_60x300_d64_w9": (60, 300, 64, 9, 21, {}),
    "bm_61x300_d64_w21": (61, 300, 64, 21, 22, {}),
    "bm_64x260_d32_w11_minneg": (64, 260, 32, 11, 23, dict(minDisparity=-16)),
    "bm_64x260_d48_w15_nouniq": (64, 260, 48, 15, 24, dict(uniquenessRatio=0, textureThreshold=0)),
    "bm_64x260_d48_w15_speckle": (64, 260, 48, 15, 25, dict(speckleWindowSize=50, speckleRange=16, preFilterCap=63)),
    "bm_kitti_d128_w27": (375, 1242, 128, 27, 0, {}),
}


def cv_bm(l, r, p):
    """Reconstruct3D.cpp:378-380 + :45-46: StereoBM::create(16, 21), setBlockSize, setNumDisparities."""
    m = cv2.StereoBM_create(16, 21)
    m.setBlockSize(p["blockSize"]); m.setNumDisparities(p["numDisparities"])
    m.setMinDisparity(p["minDisparity"]); m.setPreFilterCap(p["preFilterCap"]); m.setTextureThreshold(p["textureThreshold"])
    m.setUniquenessRatio(p["uniquenessRatio"]); m.setSpeckleWindowSize(p["speckleWindowSize"]); m.setSpeckleRange(p["speckleRange"])
    return m.compute(l, r)


def bm_cases():
    for name, (H, W, D, ws, seed, kw) in BM_CASES.items():
        lb, rb = O.synth_pair(H, W, D, seed, color=True)
        gl = cv2.cvtColor(lb, cv2.COLOR_BGR2GRAY); gr = cv2.cvtColor(rb, cv2.COLOR_BGR2GRAY)
        p = O.reference_bm_params(ws, D); p.update(kw)
        disp = cv_bm(gl, gr, p)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), disp=disp, meta=np.array([H, W, D, ws, seed], np.int64),
                            params=np.array([p[k] for k in sorted(p)], np.int64), param_names=np.array(sorted(p)),
                            crc=np.array([zlib.crc32(lb.tobytes()), zlib.crc32(rb.tobytes())], np.uint64))
        print(name, "valid %.1f%%" % (100 * (disp >= p["minDisparity"] * 16).mean()))


# ---- stereo rectification of the inputs: Reconstruct3D::RectifyImages (Reconstruct3D.cpp:330-361) ----
def _rot(rx, ry, rz):
    R, _ = cv2.Rodrigues(np.array([rx, ry, rz], np.float64))
    return R


_K_KITTI = np.array([[959.791, 0, 696.0217], [0, 956.9251, 224.1806], [0, 0, 1]], np.float64)   # calib_kitti_rectified.json, left
RECT_CASES = {   # name: (H, W, channels, K, D, R, P, image seed, keep full arrays)
    "rect_small_k5": (94, 311, 3, _K_KITTI / np.array([[4.0], [4.0], [1.0]]),
                      [-0.3728755, 0.2037299, 0.002219027, 0.001383707, -0.07233722], _rot(0.009, -0.004, 0.012),
                      np.array([[180.4, 0, 152.4, 0], [0, 180.4, 43.2, 0], [0, 0, 1, 0]], np.float64), 31, True),
    "rect_small_k8_border": (94, 311, 3, _K_KITTI / np.array([[4.0], [4.0], [1.0]]),
                             [-0.28, 0.09, 0.0011, -0.0007, 0.01, 0.02, -0.01, 0.003], _rot(-0.05, 0.08, -0.03),
                             np.array([[225.0, 0, 162.5, -100.0], [0, 225.0, 50.0, 0], [0, 0, 1, 0]], np.float64), 32, True),
    "rect_small_nodist_shift": (80, 300, 1, np.array([[300.0, 0, 150.0], [0, 300.0, 40.0], [0, 0, 1]]), [], np.eye(3),
                                np.array([[300.0, 0, 143.3], [0, 300.0, 44.7], [0, 0, 1]], np.float64), 33, True),
    "rect_small_k12": (90, 280, 4, np.array([[240.0, 0, 140.0], [0, 243.0, 45.0], [0, 0, 1]]),
                       [-0.25, 0.07, 0.0013, -0.0004, 0.012, 0.018, -0.012, 0.002, 0.001, -0.0005, 0.0007, 0.0002], _rot(0.03, 0.02, 0.05),
                       np.array([[200.0, 0, 150.0], [0, 202.5, 47.5], [0, 0, 1]], np.float64), 34, True),
    "rect_kitti_k5": (375, 1242, 3, _K_KITTI, [-0.3728755, 0.2037299, 0.002219027, 0.001383707, -0.07233722], _rot(0.009, -0.004, 0.012),
                      np.array([[721.5377, 0, 609.5593, 0], [0, 721.5377, 172.854, 0], [0, 0, 1, 0]], np.float64), 35, False),
}


def rect_image(H, W, ch, seed):
    img = np.random.default_rng(seed).integers(0, 256, (H, W, ch), dtype=np.uint8)
    return img[:, :, 0].copy() if ch == 1 else img


def rect_cases():
    for name, (H, W, ch, K, D, R, P, seed, full) in RECT_CASES.items():
        D = np.asarray(D, np.float64)
        m1, m2 = cv2.initUndistortRectifyMap(K, D, R, P, (W, H), cv2.CV_16SC2)
        img = rect_image(H, W, ch, seed)
        out = cv2.remap(img, m1, m2, cv2.INTER_LINEAR)
        d = dict(K=K, D=D, R=R, P=P, meta=np.array([H, W, ch, seed], np.int64),
                 crc=np.array([zlib.crc32(img.tobytes()), zlib.crc32(m1.tobytes()), zlib.crc32(m2.tobytes()), zlib.crc32(out.tobytes())], np.uint64))
        if full:
            d.update(map1=m1, map2=m2, out=out)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
        inside = ((m1[..., 0] >= 0) & (m1[..., 0] < W - 1) & (m1[..., 1] >= 0) & (m1[..., 1] < H - 1)).mean()
        print(name, "interior taps %.1f%%" % (100 * inside))


# ---- the full-size BASELINE.json shapes (configs[2], configs[3]) at the reference's block size: too large to store whole,
# so a fixture keeps the CRC32 of cv2's disparity / pruned disparity / mask plus every 16th row of the disparity
BIG_CASES = {   # name: (H, W, D, block_size, mode, seed)
    "hd_d256_bs27": (1080, 1920, 256, 27, 0, 5),
    "hd_d256_bs27_hh": (1080, 1920, 256, 27, 1, 5),
    "hd_d256_bs5": (1080, 1920, 256, 5, 0, 6),
    "spi_d192_bs27_hh": (720, 1280, 192, 27, 1, 5),
}


def big_cases():
    for name, (H, W, D, bs, mode, seed) in BIG_CASES.items():
        lb, rb = O.synth_pair(H, W, D, seed, color=True)
        gl = cv2.cvtColor(lb, cv2.COLOR_BGR2GRAY); gr = cv2.cvtColor(rb, cv2.COLOR_BGR2GRAY)
        p = O.reference_sgbm_params(block_size=bs, num_disparities=D, mode=mode)
        disp = cv_sgbm(gl, gr, p, True)
        dn = cv2.normalize(disp, None, 0, 255, cv2.NORM_MINMAX, cv2.CV_8U)
        _, std = cv2.meanStdDev(dn)
        _, mask = cv2.threshold(dn, std[0][0] * 1.3, 255, cv2.THRESH_BINARY)
        pruned = disp.copy(); pruned[mask == 0] = 0
        dispf = cv2.multiply(pruned.astype(np.float32), 1.0 / 16.0)
        mask2 = mask.copy(); mask2[dispf <= 0] = 0
        np.savez_compressed(os.path.join(HERE, name + ".npz"), meta=np.array([H, W, D, bs, mode, 0, seed], np.int64),
                            crc=np.array([zlib.crc32(lb.tobytes()), zlib.crc32(rb.tobytes())], np.uint64),
                            disp_crc=np.array([zlib.crc32(disp.tobytes())], np.uint64), rows16=disp[::16].copy(),
                            valid_count=np.array([int((disp >= 0).sum())], np.int64),
                            pruned_crc=np.array([zlib.crc32(pruned.tobytes())], np.uint64), mask_crc=np.array([zlib.crc32(mask2.tobytes())], np.uint64),
                            sigma=np.float64(std[0][0]), cv_version=np.array(cv2.__version__))
        print(name, "valid %.1f%%" % (100 * (disp >= 0).mean()), os.path.getsize(os.path.join(HERE, name + ".npz")) // 1024, "KiB")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "big":
        big_cases()
        sys.exit(0)
    big_cases()
    rect_cases()
    bm_cases()
    special()
    main()
