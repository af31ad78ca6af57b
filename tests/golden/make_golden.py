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


if __name__ == "__main__":
    bm_cases()
    special()
    main()
