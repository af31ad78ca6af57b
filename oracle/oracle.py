"""ctypes binding of the CPU ORACLE (oracle/cvr_oracle.c) -- test infrastructure only.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product library (libcvrecon_b200.so) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libcvr_oracle.so")

MODE_SGBM = 0
MODE_HH = 1


class SgbmParams(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
        "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")]


POINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("b", "u1"), ("g", "u1"), ("r", "u1"), ("a", "u1")])


def build(force=False):
    src = os.path.join(_HERE, "cvr_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_sgbm_compute.restype = C.c_int
        _lib.orc_generate_point_cloud.restype = C.c_int
        _lib.orc_triangulate.restype = C.c_int
        _lib.orc_voxel_grid.restype = C.c_int
        _lib.orc_sor.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def reference_sgbm_params(block_size=27, num_disparities=128, uniqueness_ratio=10, speckle_range=2,
                          speckle_window_size=100, pre_filter_cap=10, min_disparity=0, mode=MODE_SGBM,
                          disp12_max_diff=0):
    """Parameter mapping of Reconstruct3D::ConfigureSteoreoMatcher (Reconstruct3D.cpp:54-64):
    P1 = 8*3*bs^2, P2 = 32*3*bs^2, disp12MaxDiff untouched (0), mode untouched (MODE_SGBM)."""
    return dict(minDisparity=int(min_disparity), numDisparities=int(num_disparities), blockSize=int(block_size),
                P1=8 * 3 * block_size * block_size, P2=32 * 3 * block_size * block_size,
                disp12MaxDiff=disp12_max_diff, preFilterCap=pre_filter_cap, uniquenessRatio=uniqueness_ratio,
                speckleWindowSize=speckle_window_size, speckleRange=speckle_range, mode=mode)


def gray_bgr(bgr):
    bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
    H, W, ch = bgr.shape
    out = np.empty((H, W), np.uint8)
    lib().orc_gray_bgr(_p(bgr), C.c_size_t(W * ch), W, H, ch, _p(out), C.c_size_t(W))
    return out


def sgbm(left, right, params, stages=False):
    """left/right: uint8 (H, W).  Returns disp int16 (H, W); with stages=True also a dict with
    raw (pre-median), C, S volumes and (minL, minC) stats."""
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    H, W = left.shape
    prm = SgbmParams(**params)
    disp = np.empty((H, W), np.int16)
    stats = (C.c_int * 2)()
    if not stages:
        lib().orc_sgbm_compute(_p(left), C.c_size_t(W), _p(right), C.c_size_t(W), W, H, C.byref(prm),
                               _p(disp), None, None, None, stats)
        return disp
    minD, D = params["minDisparity"], params["numDisparities"]
    W1 = (W + min(minD, 0)) - max(minD + D, 0)
    raw = np.empty((H, W), np.int16)
    Cv = np.zeros((H, max(W1, 0), D), np.int16)
    Sv = np.zeros((H, max(W1, 0), D), np.int16)
    lib().orc_sgbm_compute(_p(left), C.c_size_t(W), _p(right), C.c_size_t(W), W, H, C.byref(prm),
                           _p(disp), _p(raw), _p(Cv), _p(Sv), stats)
    return disp, dict(raw=raw, C=Cv, S=Sv, minL=stats[0], minC=stats[1])


class BmParams(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("minDisparity", "numDisparities", "blockSize", "preFilterCap", "textureThreshold",
                                      "uniquenessRatio", "speckleWindowSize", "speckleRange")]


def reference_bm_params(window_size=27, num_disparities=128):
    """Reconstruct3D::ConfigureSteoreoMatcher, STEREO_BLOCK_MATCHER branch (Reconstruct3D.cpp:43-48, :378-380):
    cv::StereoBM::create(16, 21) then setBlockSize / setNumDisparities; everything else stays at OpenCV's defaults."""
    return dict(minDisparity=0, numDisparities=int(num_disparities), blockSize=int(window_size), preFilterCap=31, textureThreshold=10,
                uniquenessRatio=15, speckleWindowSize=0, speckleRange=0)


def stereobm(left, right, params):
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    H, W = left.shape
    prm = BmParams(**params)
    disp = np.empty((H, W), np.int16)
    lib().orc_stereobm_compute(_p(left), C.c_size_t(W), _p(right), C.c_size_t(W), W, H, C.byref(prm), _p(disp))
    return disp


def median3(img):
    img = np.ascontiguousarray(img, np.int16)
    out = np.empty_like(img)
    lib().orc_median3_s16(_p(img), _p(out), img.shape[1], img.shape[0])
    return out


def filter_speckles(img, new_val, max_size, max_diff):
    out = np.ascontiguousarray(img, np.int16).copy()
    lib().orc_filter_speckles(_p(out), out.shape[1], out.shape[0], int(new_val), int(max_size), int(max_diff))
    return out


def prune(disp):
    """TrackingFrame::SetupFrame.  Returns (pruned int16, dispf f32, mask u8, norm u8, stats[mn,mx,sigma,ithr])."""
    d = np.ascontiguousarray(disp, np.int16).copy()
    H, W = d.shape
    f = np.empty((H, W), np.float32)
    m = np.empty((H, W), np.uint8)
    u8 = np.empty((H, W), np.uint8)
    st = np.zeros(4, np.float64)
    lib().orc_prune(_p(d), W, H, _p(f), _p(m), _p(u8), _p(st))
    return d, f, m, u8, st


def reproject(disp, Q, handle_missing=True):
    d = np.ascontiguousarray(disp, np.int16)
    H, W = d.shape
    Q = np.ascontiguousarray(Q, np.float32)
    xyz = np.empty((H, W, 3), np.float32)
    lib().orc_reproject(_p(d), W, H, _p(Q), int(bool(handle_missing)), _p(xyz))
    return xyz


def generate_point_cloud(disp, bgr, Q):
    d = np.ascontiguousarray(disp, np.int16)
    bgr = np.ascontiguousarray(bgr, np.uint8)
    H, W = d.shape
    Q = np.ascontiguousarray(Q, np.float32)
    out = np.empty(H * W, POINT_DTYPE)
    n = lib().orc_generate_point_cloud(_p(d), _p(bgr), C.c_size_t(W * 3), W, H, _p(Q), _p(out))
    return out[:n].copy()


def triangulate(dispf, bgr, mask, f, b, cx, cy):
    d = np.ascontiguousarray(dispf, np.float32)
    bgr = np.ascontiguousarray(bgr, np.uint8)
    H, W = d.shape
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.uint8)
    out = np.empty(H * W, POINT_DTYPE)
    n = lib().orc_triangulate(_p(d), _p(bgr), C.c_size_t(W * 3), _p(mask), W, H,
                              C.c_float(f), C.c_float(b), C.c_float(cx), C.c_float(cy), _p(out))
    return out[:n].copy()


def voxel_grid(pts, leaf):
    pts = np.ascontiguousarray(pts, POINT_DTYPE)
    n = len(pts)
    out = np.empty(max(n, 1), POINT_DTYPE)
    counts = np.zeros(max(n, 1), np.int32)
    pt = C.c_int(0)
    lx, ly, lz = (leaf, leaf, leaf) if np.isscalar(leaf) else leaf
    m = lib().orc_voxel_grid(_p(pts), n, C.c_float(lx), C.c_float(ly), C.c_float(lz), _p(out), _p(counts), C.byref(pt))
    return out[:m].copy(), counts[:m].copy(), bool(pt.value)


def radius_outlier_removal(pts, radius, min_pts):
    pts = np.ascontiguousarray(pts, POINT_DTYPE)
    n = len(pts)
    out = np.empty(max(n, 1), POINT_DTYPE)
    keep = np.zeros(max(n, 1), np.uint8)
    lib().orc_radius_outlier_removal.restype = C.c_int
    m = lib().orc_radius_outlier_removal(_p(pts), n, C.c_float(radius), int(min_pts), _p(out), _p(keep))
    return out[:m].copy(), keep[:n].copy()


def transform_cloud(pts, T):
    pts = np.ascontiguousarray(pts, POINT_DTYPE)
    T = np.ascontiguousarray(T, np.float32)
    out = np.empty(len(pts), POINT_DTYPE)
    lib().orc_transform_cloud(_p(pts), len(pts), _p(T), _p(out))
    return out


def sor(pts, k, alpha):
    """Brute-force exact kNN statistical outlier removal.  Returns (kept points, mean dists, thr, keep mask)."""
    pts = np.ascontiguousarray(pts, POINT_DTYPE)
    n = len(pts)
    out = np.empty(max(n, 1), POINT_DTYPE)
    dist = np.zeros(max(n, 1), np.float32)
    keep = np.zeros(max(n, 1), np.uint8)
    thr = C.c_double(0)
    m = lib().orc_sor(_p(pts), n, int(k), C.c_double(alpha), _p(out), _p(dist), C.byref(thr), _p(keep))
    return out[:m].copy(), dist[:n].copy(), thr.value, keep[:n].copy()


def rectify_inverse(R, P):
    """inverse(P[:, :3] @ R) as initUndistortRectifyMap computes it (3x3 closed form in double)."""
    R = np.ascontiguousarray(R, np.float64).reshape(3, 3)
    P = np.ascontiguousarray(P, np.float64)
    P = P.reshape(3, -1)
    ir = np.empty(9, np.float64)
    lib().orc_rectify_inverse(_p(R), _p(P), int(P.shape[1]), _p(ir))
    return ir


def init_undistort_rectify_map(K, D, R, P, W, H):
    """cv::initUndistortRectifyMap(K, D, R, P, (W, H), CV_16SC2): returns (map1 int16 (H, W, 2), map2 uint16 (H, W))."""
    K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
    D = np.ascontiguousarray(D, np.float64).ravel()
    assert D.size <= 12, "the tilt terms (tauX, tauY) are not restated"
    R = np.ascontiguousarray(R, np.float64).reshape(3, 3)
    P = np.ascontiguousarray(P, np.float64).reshape(3, -1)
    m1 = np.empty((H, W, 2), np.int16)
    m2 = np.empty((H, W), np.uint16)
    lib().orc_init_undistort_rectify_map(_p(K), _p(D) if D.size else None, int(D.size), _p(R), _p(P), int(P.shape[1]), int(W), int(H), _p(m1), _p(m2))
    return m1, m2


def remap_bilinear(src, map1, map2):
    """cv::remap(src, map1, map2, INTER_LINEAR) for 8-bit images with CV_16SC2 + CV_16UC1 maps, constant border 0."""
    src = np.ascontiguousarray(src, np.uint8)
    sh, sw = src.shape[:2]
    cn = 1 if src.ndim == 2 else src.shape[2]
    map1 = np.ascontiguousarray(map1, np.int16)
    map2 = np.ascontiguousarray(map2, np.uint16)
    H, W = map2.shape
    dst = np.empty((H, W) if src.ndim == 2 else (H, W, cn), np.uint8)
    lib().orc_remap_bilinear_u8(_p(src), C.c_size_t(sw * cn), sw, sh, cn, _p(map1), _p(map2), W, H, _p(dst), C.c_size_t(W * cn))
    return dst


# --------------------------------------------------------------------------------------------
# Synthetic rectified pair generator: lives in the package (numpy only, no compute-path code) so
# that bench.py can synthesise inputs without importing the oracle; re-exported here for the tests.
import importlib as _il
import sys as _sys

if os.path.dirname(_HERE) not in _sys.path:
    _sys.path.insert(0, os.path.dirname(_HERE))
_synth = _il.import_module("master-thesis_b200.synth")
synth_pair = _synth.synth_pair
satdrift_pair = _synth.satdrift_pair


def icp(source, target, max_dist=5.0, max_iter=50, trans_eps=1e-8, fit_eps=1.0):
    """pcl::IterativeClosestPoint::align with the reference's settings (MappingSystem.cpp:321-341); brute-force neighbours.
    Returns (T 4x4 f32, aligned points, iterations, fitness score, state, correspondences of the last iteration)."""
    src = np.ascontiguousarray(source, POINT_DTYPE)
    tgt = np.ascontiguousarray(target, POINT_DTYPE)
    T = np.zeros(16, np.float32)
    out = np.empty(max(len(src), 1), POINT_DTYPE)
    it, nc = C.c_int(0), C.c_int(0)
    fit = C.c_double(0)
    L = lib()
    L.orc_icp.restype = C.c_int
    st = L.orc_icp(_p(src), len(src), _p(tgt), len(tgt), C.c_double(max_dist), int(max_iter), C.c_double(trans_eps), C.c_double(fit_eps),
                   _p(T), _p(out), C.byref(it), C.byref(fit), C.byref(nc))
    return T.reshape(4, 4).copy(), out[:len(src)].copy(), it.value, fit.value, st, nc.value


def umeyama_from_sums(sums):
    s = np.ascontiguousarray(sums, np.float64)
    T = np.zeros(16, np.float32)
    lib().orc_umeyama_from_sums(_p(s), _p(T))
    return T.reshape(4, 4).copy()
