"""master-thesis_b200 -- ctypes binding of libcvrecon_b200.so (include/cvrecon_b200.h).

The product is the CUDA library; this module is the thin host-side harness used by tests/ and
bench.py.  Method names mirror the reference's C++ interface for this path
(cv_reconstruct/include/reconstruct/Reconstruct3D.hpp:25-103, TrackingFrame.cpp:25-67,
PointCloudPostProcessor.hpp:38) so that parity tests read like calls into the reference.

There is NO CPU fallback: if the library is missing or no B200-class GPU is present, construction
raises.  Nothing in this package imports oracle/.

Import with:  importlib.import_module("master-thesis_b200")  (the directory name has a hyphen).
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libcvrecon_b200.so")

MODE_SGBM, MODE_HH = 0, 1
MEM_HOST, MEM_DEVICE = 0, 1
CLOUD_TRIANGULATE, CLOUD_REPROJECT = 0, 1
MATCHER_SGBM, MATCHER_BM = 0, 1
STAGE_DISPARITY, STAGE_CLOUD, STAGE_VOXEL, STAGE_SOR = 1, 2, 4, 8

POINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("b", "u1"), ("g", "u1"), ("r", "u1"), ("a", "u1")])

# every symbol include/cvrecon_b200.h declares
EXPORTS = [
    "cvr_default_params", "cvr_create", "cvr_destroy", "cvr_last_error", "cvr_version", "cvr_disparity", "cvr_prune",
    "cvr_reproject_cloud", "cvr_triangulate", "cvr_voxel_grid", "cvr_sor", "cvr_process_batch", "cvr_points_to_pcl", "cvr_save_pcd",
    "cvr_debug_fetch", "cvr_launch_count", "cvr_last_timings", "cvr_set_profiling", "cvr_prune_noalloc",
    "cvr_submit_batch", "cvr_wait_batch", "cvr_radius_outlier_removal", "cvr_transform_cloud",
    "cvr_set_rectification", "cvr_rectify", "cvr_get_rectification_maps", "cvr_suggest_batch",
    "cvr_comm_get_unique_id", "cvr_comm_init", "cvr_comm_attach", "cvr_comm_destroy", "cvr_gather_clouds", "cvr_gather_layout",
    "cvr_filter_cloud", "cvr_icp_default_params", "cvr_icp_align", "cvr_decode_png_batch",
]
COMM_ID_BYTES = 128


class SgbmParams(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "min_disparity", "num_disparities", "block_size", "P1", "P2", "disp12_max_diff", "pre_filter_cap",
        "uniqueness_ratio", "speckle_window_size", "speckle_range", "mode")]


class Params(C.Structure):
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("sgbm", SgbmParams), ("Q", C.c_float * 16),
        ("tri_f", C.c_float), ("tri_b", C.c_float), ("tri_cx", C.c_float), ("tri_cy", C.c_float),
        ("prune_sigma_factor", C.c_double), ("voxel_leaf", C.c_float * 3), ("sor_mean_k", C.c_int),
        ("sor_stddev_mul", C.c_double), ("max_batch", C.c_int), ("cloud_mode", C.c_int), ("max_points", C.c_int),
        ("filter_only", C.c_int), ("block_matcher", C.c_int), ("bm_window_size", C.c_int), ("bm_num_disparities", C.c_int),
        ("bm_min_disparity", C.c_int), ("bm_pre_filter_cap", C.c_int), ("bm_texture_threshold", C.c_int), ("bm_uniqueness_ratio", C.c_int),
        ("bm_speckle_window_size", C.c_int), ("bm_speckle_range", C.c_int), ("requires_rectification", C.c_int),
    ]


class Batch(C.Structure):
    _fields_ = [
        ("mem", C.c_int), ("channels", C.c_int),
        ("left", C.c_void_p), ("left_step", C.c_size_t), ("left_frame_stride", C.c_size_t),
        ("right", C.c_void_p), ("right_step", C.c_size_t), ("right_frame_stride", C.c_size_t),
        ("disp", C.c_void_p), ("disp_step", C.c_size_t), ("disp_frame_stride", C.c_size_t),
        ("cloud", C.c_void_p), ("cloud_capacity", C.c_size_t), ("cloud_count", C.c_void_p), ("stages", C.c_int),
        ("cloud_packed", C.c_void_p), ("cloud_packed_capacity", C.c_size_t), ("cloud_offsets", C.c_void_p),
    ]


class IcpParams(C.Structure):
    _fields_ = [("max_correspondence_distance", C.c_double), ("max_iterations", C.c_int), ("transformation_epsilon", C.c_double),
                ("euclidean_fitness_epsilon", C.c_double)]


class IcpResult(C.Structure):
    _fields_ = [("T", C.c_float * 16), ("iterations", C.c_int), ("converged", C.c_int), ("state", C.c_int), ("correspondences", C.c_int),
                ("fitness", C.c_double)]


class CvrError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("cvrecon_b200 error %d: %s" % (code, msg))
        self.code = code


def build(force=False):
    """Compile the CUDA library in-tree (nvcc, sm_100a)."""
    srcdir = os.path.join(_HERE, "csrc")
    cmd = ["make", "-C", srcdir, "-j8", "-s"]
    if force:
        subprocess.check_call(["make", "-C", srcdir, "clean", "-s"])
    subprocess.check_call(cmd)
    return LIB_PATH


_lib = None


def load_library():
    """dlopen the product library; raises if it is not built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CvrError(-4, "libcvrecon_b200.so is not built (run __graft_entry__.build()); there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    lib.cvr_last_error.restype = C.c_char_p
    lib.cvr_last_error.argtypes = [C.c_void_p]
    lib.cvr_version.restype = C.c_char_p
    lib.cvr_launch_count.restype = C.c_longlong
    lib.cvr_launch_count.argtypes = [C.c_void_p]
    lib.cvr_create.argtypes = [C.POINTER(Params), C.c_int, C.POINTER(C.c_void_p)]
    lib.cvr_destroy.argtypes = [C.c_void_p]
    lib.cvr_default_params.argtypes = [C.POINTER(Params), C.c_int, C.c_int]
    lib.cvr_default_params.restype = None
    lib.cvr_disparity.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t]
    lib.cvr_prune.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    lib.cvr_reproject_cloud.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    lib.cvr_triangulate.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    lib.cvr_voxel_grid.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.cvr_sor.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    lib.cvr_process_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(Batch)]
    lib.cvr_radius_outlier_removal.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    lib.cvr_transform_cloud.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.cvr_set_rectification.argtypes = [C.c_void_p] + [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int] * 2
    lib.cvr_rectify.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    lib.cvr_get_rectification_maps.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.cvr_suggest_batch.argtypes = [C.c_void_p, C.c_int]
    lib.cvr_submit_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(Batch)]
    lib.cvr_wait_batch.argtypes = [C.c_void_p]
    lib.cvr_debug_fetch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.cvr_last_timings.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_float), C.c_int]
    lib.cvr_set_profiling.argtypes = [C.c_void_p, C.c_int]
    lib.cvr_points_to_pcl.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.cvr_points_to_pcl.restype = None
    lib.cvr_save_pcd.argtypes = [C.c_char_p, C.c_void_p, C.c_int]
    lib.cvr_save_pcd.restype = C.c_int
    lib.cvr_prune_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.cvr_filter_cloud.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    lib.cvr_icp_default_params.argtypes = [C.POINTER(IcpParams)]
    lib.cvr_icp_default_params.restype = None
    lib.cvr_icp_align.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(IcpParams), C.c_void_p, C.POINTER(IcpResult)]
    lib.cvr_decode_png_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_int, C.c_void_p, C.c_size_t, C.c_size_t]
    lib.cvr_comm_get_unique_id.argtypes = [C.c_void_p]
    lib.cvr_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.cvr_comm_attach.argtypes = [C.c_void_p, C.c_void_p]
    lib.cvr_comm_destroy.argtypes = [C.c_void_p]
    lib.cvr_gather_clouds.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_void_p]
    lib.cvr_gather_layout.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    _lib = lib
    return lib


def comm_unique_id():
    """ncclGetUniqueId through the library (rank 0 calls it and hands the 128 bytes to the other ranks)."""
    buf = (C.c_ubyte * COMM_ID_BYTES)()
    rc = load_library().cvr_comm_get_unique_id(buf)
    if rc != 0:
        raise CvrError(rc, load_library().cvr_last_error(None).decode())
    return bytes(buf)


def gather_layout(headers):
    """cvr_gather_layout: headers = int32 array [nranks][2 + max_batch] of {nframes, total, counts...}.
    Returns (rank_offset[nranks+1] int64, frame_offsets[total_frames+1] int32, frames_of_rank[nranks] int32)."""
    h = np.ascontiguousarray(headers, np.int32)
    nr, hi = h.shape
    roff = np.zeros(nr + 1, np.int64)
    foff = np.zeros(int(h[:, 0].clip(0).sum()) + 1, np.int32)
    fr = np.zeros(nr, np.int32)
    rc = load_library().cvr_gather_layout(_ptr(h), nr, hi, _ptr(roff), _ptr(foff), len(foff), _ptr(fr))
    if rc < 0:
        raise CvrError(rc, "cvr_gather_layout failed")
    return roff, foff[:rc + 1], fr


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


# ---------------------------------------------------------------------------------------------
# Config: the same JSON keys the reference parses (cv_reconstruct/src/config/ConfigParser.cpp:18-67)
# plus the point_cloud_post_processing keys that are present in config_default.json but unparsed.
DEFAULT_CONFIG = {
    "config": {
        "reconstruction": {
            "requires_rectification": False,
            "block_matcher": "stereo_sgbm",
            "SBM": {"window_size": 27, "num_disparities": 128},
            "SGBM": {"block_size": 27, "num_disparities": 128, "uniqueness_ratio": 10, "speckle_range": 2,
                     "speckle_window_size": 100, "pre_filter_cap": 10, "min_disparity": 0.0},
        },
        "point_cloud_post_processing": {"outlier_min_k": 30, "outlier_std_threshold": 0.5},
    }
}

KITTI_CALIB = dict(fx_l=959.791, fy_l=956.9251, cx_l=696.0217, cy_l=224.1806, fx_r=903.7596, T0=0.532671)


def parse_config(path_or_dict=None):
    cfg = DEFAULT_CONFIG if path_or_dict is None else path_or_dict
    if isinstance(cfg, (str, os.PathLike)):
        with open(cfg) as f:
            cfg = json.load(f)
    rec = cfg["config"]["reconstruction"]
    # ConfigParser.cpp:40-50: "stereo_bm" / "stereo_sgbm", anything else falls back to the block matcher
    bm = rec.get("block_matcher", "stereo_bm") != "stereo_sgbm"
    s = rec["SGBM"]
    sbm = rec.get("SBM", {"window_size": 21, "num_disparities": 16})
    pp = cfg["config"].get("point_cloud_post_processing", {})
    return dict(requires_rectification=bool(rec.get("requires_rectification", False)),       # ConfigParser.cpp:39
                block_matcher=MATCHER_BM if bm else MATCHER_SGBM, window_size=int(sbm["window_size"]),
                bm_num_disparities=int(sbm["num_disparities"]), block_size=int(s["block_size"]), num_disparities=int(s["num_disparities"]),
                uniqueness_ratio=int(s["uniqueness_ratio"]), speckle_range=int(s["speckle_range"]),
                speckle_window_size=int(s["speckle_window_size"]), pre_filter_cap=int(s["pre_filter_cap"]),
                min_disparity=int(s["min_disparity"]), outlier_min_k=int(pp.get("outlier_min_k", 30)),
                outlier_std_threshold=float(pp.get("outlier_std_threshold", 0.5)))


class Reconstruct3D:
    """Mirror of Reconstruct::Reconstruct3D (Reconstruct3D.hpp:25-103) on top of the C ABI.

    Reconstruct3D(width, height, config=..., calib=...) does what the reference constructor does
    (Reconstruct3D.cpp:26-69): maps the SGBM config (P1 = 8*3*bs^2, P2 = 32*3*bs^2), builds Q."""

    def __init__(self, width, height, config=None, calib=None, mode=MODE_SGBM, disp12_max_diff=0, P1=None, P2=None,
                 max_batch=1, cloud_mode=CLOUD_TRIANGULATE, voxel_leaf=0.2, device=0, max_points=0, filter_only=False, Q=None):
        self.lib = load_library()
        cfg = parse_config(config) if not (isinstance(config, dict) and "block_size" in config) else config
        cal = dict(KITTI_CALIB)
        if calib:
            cal.update(calib)
        p = Params()
        self.lib.cvr_default_params(C.byref(p), int(width), int(height))
        bs = cfg["block_size"]
        p.sgbm.block_size = bs
        p.sgbm.num_disparities = cfg["num_disparities"]
        p.sgbm.uniqueness_ratio = cfg["uniqueness_ratio"]
        p.sgbm.speckle_range = cfg["speckle_range"]
        p.sgbm.speckle_window_size = cfg["speckle_window_size"]
        p.sgbm.pre_filter_cap = cfg["pre_filter_cap"]
        p.sgbm.min_disparity = cfg["min_disparity"]
        p.sgbm.P1 = 8 * 3 * bs * bs if P1 is None else P1          # Reconstruct3D.cpp:63
        p.sgbm.P2 = 32 * 3 * bs * bs if P2 is None else P2         # Reconstruct3D.cpp:64
        p.sgbm.disp12_max_diff = disp12_max_diff
        p.sgbm.mode = mode
        if Q is None:
            f = np.float32((np.float32(cal["fx_l"]) + np.float32(cal["fx_r"])) / np.float32(2.0))   # Reconstruct3D.cpp:33
            Qm = np.eye(4, dtype=np.float32)
            Qm[2, 2] = np.float32(0.005 * float(f))                                                   # Reconstruct3D.cpp:34
        else:
            Qm = np.asarray(Q, np.float32).reshape(4, 4)
        for i, v in enumerate(Qm.reshape(-1)):
            p.Q[i] = float(v)
        p.tri_f = float(np.float32((np.float32(cal["fx_l"]) + np.float32(cal["fy_l"])) / np.float32(2.0)))  # Reconstruct3D.cpp:168
        p.tri_b = float(np.float32(cal["T0"]))
        p.tri_cx = float(np.float32(cal["cx_l"]))
        p.tri_cy = float(np.float32(cal["cy_l"]))
        leaf = (voxel_leaf,) * 3 if np.isscalar(voxel_leaf) else voxel_leaf
        for i in range(3):
            p.voxel_leaf[i] = float(leaf[i])
        p.sor_mean_k = cfg.get("outlier_min_k", 30)
        p.sor_stddev_mul = cfg.get("outlier_std_threshold", 0.5)
        p.max_batch = int(max_batch)
        p.cloud_mode = int(cloud_mode)
        p.max_points = int(max_points)
        p.block_matcher = int(cfg.get("block_matcher", MATCHER_SGBM))
        p.bm_window_size = int(cfg.get("window_size", 27))               # Reconstruct3D.cpp:45
        p.bm_num_disparities = int(cfg.get("bm_num_disparities", 128))     # Reconstruct3D.cpp:46
        for k_ in ("bm_min_disparity", "bm_pre_filter_cap", "bm_texture_threshold", "bm_uniqueness_ratio", "bm_speckle_window_size", "bm_speckle_range"):
            if k_ in cfg:
                setattr(p, k_, int(cfg[k_]))
        p.filter_only = int(bool(filter_only))
        p.requires_rectification = int(bool(cfg.get("requires_rectification", False)))
        self.params = p
        self.Q = Qm
        self.W, self.H = int(width), int(height)
        self.ctx = C.c_void_p()
        rc = self.lib.cvr_create(C.byref(p), int(device), C.byref(self.ctx))
        if rc != 0:
            raise CvrError(rc, self.lib.cvr_last_error(None).decode())
        self.max_points = max_points if max_points > 0 else self.W * self.H

    # -- plumbing --
    def _check(self, rc):
        if rc != 0:
            raise CvrError(rc, self.lib.cvr_last_error(self.ctx).decode())

    def close(self):
        if getattr(self, "ctx", None) and self.ctx.value:
            self.lib.cvr_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def suggest_batch(self, around):
        """Batch size near `around` (<= max_batch) that fills the vertical path kernel's cluster waves on this GPU."""
        n = int(self.lib.cvr_suggest_batch(self.ctx, int(around)))
        if n < 1:
            raise CvrError(n, self.lib.cvr_last_error(self.ctx).decode())
        return n

    def launch_count(self):
        return int(self.lib.cvr_launch_count(self.ctx))

    def set_profiling(self, on):
        self.lib.cvr_set_profiling(self.ctx, int(bool(on)))

    def last_timings(self):
        names = (C.c_char_p * 64)()
        ms = (C.c_float * 64)()
        n = self.lib.cvr_last_timings(self.ctx, names, ms, 64)
        return [(names[i].decode(), float(ms[i])) for i in range(n)]

    # -- Reconstruct3D::RectifyImages (Reconstruct3D.cpp:330-361) --
    def SetRectification(self, K1, D1, R1, P1, K2, D2, R2, P2):
        """The calibration RectifyImages reads from m_StereoCameraSetup (K, D per camera; Rectification.RL/RR/PL/PR):
        builds the CV_16SC2 maps of both cameras on the device, once."""
        def prep(K, D, R, P):
            K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
            D = np.ascontiguousarray(np.asarray(D, np.float64).ravel())
            R = np.ascontiguousarray(R, np.float64).reshape(3, 3)
            P = np.ascontiguousarray(P, np.float64)
            P = P.reshape(3, -1)
            return K, D, R, P
        a, b = prep(K1, D1, R1, P1), prep(K2, D2, R2, P2)
        self._rect_keep = (a, b)
        args = []
        for K, D, R, P in (a, b):
            args += [_ptr(K), _ptr(D) if D.size else None, int(D.size), _ptr(R), _ptr(P), int(P.shape[1])]
        self._check(self.lib.cvr_set_rectification(self.ctx, *args))

    def RectifyImages(self, leftImage, rightImage):
        l = np.ascontiguousarray(leftImage, np.uint8)
        r = np.ascontiguousarray(rightImage, np.uint8)
        ch = 1 if l.ndim == 2 else l.shape[2]
        ol, orr = np.empty_like(l), np.empty_like(r)
        step = self.W * ch
        self._check(self.lib.cvr_rectify(self.ctx, ch, _ptr(l), step, _ptr(r), step, _ptr(ol), step, _ptr(orr), step))
        return ol, orr

    def rectification_maps(self, camera):
        m1 = np.empty((self.H, self.W, 2), np.int16)
        m2 = np.empty((self.H, self.W), np.uint16)
        self._check(self.lib.cvr_get_rectification_maps(self.ctx, int(camera), _ptr(m1), _ptr(m2)))
        return m1, m2

    # -- cv::imdecode(png, IMREAD_COLOR) x n (MessageConverter.cpp:65-66) --
    def DecodePNG(self, files, device_out=None):
        """files: list of bytes objects (PNG files).  Returns an (n, H, W, 3) uint8 array, or fills device_out (a device pointer to
        n dense frames) when given."""
        n = len(files)
        bufs = [np.frombuffer(f, np.uint8) for f in files]
        ptrs = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        sizes = (C.c_size_t * n)(*[len(b) for b in bufs])
        step, stride = self.W * 3, self.W * 3 * self.H
        if device_out is not None:
            self._check(self.lib.cvr_decode_png_batch(self.ctx, n, ptrs, sizes, MEM_DEVICE, C.c_void_p(device_out), step, stride))
            return None
        out = np.empty((n, self.H, self.W, 3), np.uint8)
        self._check(self.lib.cvr_decode_png_batch(self.ctx, n, ptrs, sizes, MEM_HOST, _ptr(out), step, stride))
        return out

    # -- Reconstruct3D::GenerateDisparityMap (Reconstruct3D.cpp:72-84) --
    def GenerateDisparityMap(self, leftImage, rightImage):
        l = np.ascontiguousarray(leftImage, np.uint8)
        r = np.ascontiguousarray(rightImage, np.uint8)
        ch = 1 if l.ndim == 2 else l.shape[2]
        if l.shape[:2] != (self.H, self.W) or r.shape != l.shape:
            raise CvrError(-1, "image size does not match the context")
        disp = np.empty((self.H, self.W), np.int16)
        self._check(self.lib.cvr_disparity(self.ctx, _ptr(l), l.strides[0], _ptr(r), r.strides[0], ch, _ptr(disp), disp.strides[0]))
        return disp

    def debug_fetch(self, which):
        w1, dp = C.c_int(), C.c_int()
        self._check(self.lib.cvr_debug_fetch(self.ctx, which, None, 0, C.byref(w1), C.byref(dp)))
        if which in (0, 1):
            out = np.empty((self.H, max(w1.value, 0), dp.value), np.int16)
        else:
            out = np.empty((self.H, self.W), np.int16)
        self._check(self.lib.cvr_debug_fetch(self.ctx, which, _ptr(out), out.nbytes, C.byref(w1), C.byref(dp)))
        return out

    def sor_debug(self, n):
        """Mean k-NN distances (input order) and (sum, sq, thr, valid, kept) of the last RemoveOutliers call (test hook)."""
        d = np.empty(n, np.float32)
        self._check(self.lib.cvr_debug_fetch(self.ctx, 4, _ptr(d), d.nbytes, None, None))
        st = np.zeros(1, np.dtype([("sum", "<f8"), ("sq", "<f8"), ("thr", "<f8"), ("valid", "<i4"), ("kept", "<i4")]))
        self._check(self.lib.cvr_debug_fetch(self.ctx, 5, _ptr(st), st.nbytes, None, None))
        return d, st[0]

    # -- TrackingFrame::SetupFrame / PruneDisparityImage (TrackingFrame.cpp:25-67) --
    def SetupFrame(self, disparity):
        d = np.ascontiguousarray(disparity, np.int16).copy()
        f = np.empty((self.H, self.W), np.float32)
        m = np.empty((self.H, self.W), np.uint8)
        self._check(self.lib.cvr_prune(self.ctx, _ptr(d), d.strides[0], _ptr(f), _ptr(m)))
        return d, f, m

    # -- Reconstruct3D::GeneratePointCloud (Reconstruct3D.cpp:106-153) --
    def GeneratePointCloud(self, disparity, cameraImage):
        d = np.ascontiguousarray(disparity, np.int16)
        img = np.ascontiguousarray(cameraImage, np.uint8)
        out = np.empty(self.W * self.H, POINT_DTYPE)
        n = C.c_int()
        self._check(self.lib.cvr_reproject_cloud(self.ctx, _ptr(d), d.strides[0], _ptr(img), img.strides[0], _ptr(out), len(out), C.byref(n)))
        return out[:n.value].copy()

    # -- Reconstruct3D::Triangulate3D (Reconstruct3D.cpp:156-216) --
    def Triangulate3D(self, disparity, cameraImage, mask=None):
        d = np.ascontiguousarray(disparity, np.float32)
        img = np.ascontiguousarray(cameraImage, np.uint8)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        out = np.empty(self.W * self.H, POINT_DTYPE)
        n = C.c_int()
        self._check(self.lib.cvr_triangulate(self.ctx, _ptr(d), d.strides[0], _ptr(img), img.strides[0], _ptr(mask),
                                             0 if mask is None else mask.strides[0], _ptr(out), len(out), C.byref(n)))
        return out[:n.value].copy()

    # -- pcl::VoxelGrid::filter at MappingSystem.cpp:297-300 --
    def VoxelGrid(self, cloud):
        pts = np.ascontiguousarray(cloud, POINT_DTYPE)
        out = np.empty(max(len(pts), 1), POINT_DTYPE)
        m, pt = C.c_int(), C.c_int()
        self._check(self.lib.cvr_voxel_grid(self.ctx, _ptr(pts), len(pts), _ptr(out), len(out), C.byref(m), C.byref(pt)))
        return out[:m.value].copy(), bool(pt.value)

    # -- PointCloudPostProcessor::RemoveOutliers (PointCloudPostProcessor.cpp:38-41) --
    def RemoveOutliers(self, cloud):
        pts = np.ascontiguousarray(cloud, POINT_DTYPE)
        out = np.empty(max(len(pts), 1), POINT_DTYPE)
        k = C.c_int()
        self._check(self.lib.cvr_sor(self.ctx, _ptr(pts), len(pts), _ptr(out), len(out), C.byref(k)))
        return out[:k.value].copy()

    def FilterCloud(self, cloud, stages=STAGE_VOXEL | STAGE_SOR):
        """VoxelGrid and/or RemoveOutliers in one call (host arrays)."""
        pts = np.ascontiguousarray(cloud, POINT_DTYPE)
        out = np.empty(max(len(pts), 1), POINT_DTYPE)
        m = C.c_int()
        self._check(self.lib.cvr_filter_cloud(self.ctx, MEM_HOST, _ptr(pts), len(pts), int(stages), _ptr(out), len(out), C.byref(m)))
        return out[:m.value].copy()

    def filter_cloud_device(self, in_ptr, n, stages, out_ptr, capacity):
        """cvr_filter_cloud on device pointers; returns the output count."""
        m = C.c_int()
        self._check(self.lib.cvr_filter_cloud(self.ctx, MEM_DEVICE, C.c_void_p(in_ptr), int(n), int(stages), C.c_void_p(out_ptr), int(capacity), C.byref(m)))
        return m.value

    # -- pcl::IterativeClosestPoint::align at MappingSystem.cpp:321-341 --
    def AlignICP(self, source, target, **kw):
        """Returns (T 4x4 f32, aligned source points, IcpResult)."""
        src = np.ascontiguousarray(source, POINT_DTYPE)
        tgt = np.ascontiguousarray(target, POINT_DTYPE)
        prm = IcpParams()
        self.lib.cvr_icp_default_params(C.byref(prm))
        for k_, v in kw.items():
            setattr(prm, k_, v)
        res = IcpResult()
        out = np.empty(len(src), POINT_DTYPE)
        self._check(self.lib.cvr_icp_align(self.ctx, MEM_HOST, _ptr(src), len(src), _ptr(tgt), len(tgt), C.byref(prm), _ptr(out), C.byref(res)))
        return np.array(list(res.T), np.float32).reshape(4, 4), out, res

    # -- pcl::RadiusOutlierRemoval::filter at MappingSystem.cpp:303-307 --
    def RadiusOutlierRemoval(self, cloud, radius=1.0, min_neighbors=300):
        pts = np.ascontiguousarray(cloud, POINT_DTYPE)
        out = np.empty(max(len(pts), 1), POINT_DTYPE)
        k = C.c_int()
        self._check(self.lib.cvr_radius_outlier_removal(self.ctx, _ptr(pts), len(pts), float(radius), int(min_neighbors), _ptr(out), len(out), C.byref(k)))
        return out[:k.value].copy()

    # -- pcl::transformPointCloud at MappingSystem.cpp:318 --
    def TransformPointCloud(self, cloud, T):
        pts = np.ascontiguousarray(cloud, POINT_DTYPE)
        Tm = np.ascontiguousarray(T, np.float32).reshape(4, 4)
        out = np.empty(len(pts), POINT_DTYPE)
        self._check(self.lib.cvr_transform_cloud(self.ctx, _ptr(pts), len(pts), _ptr(Tm), _ptr(out)))
        return out

    # -- ReconstructionSystem::ProcessStereoFrame chain for a batch of frames (host buffers) --
    def ProcessFrames(self, lefts, rights, stages=0, want_disp=False, cloud_capacity=None):
        L = np.ascontiguousarray(lefts, np.uint8)
        R = np.ascontiguousarray(rights, np.uint8)
        B = L.shape[0]
        ch = 1 if L.ndim == 3 else L.shape[3]
        cap = cloud_capacity or self.W * self.H
        io = Batch()
        io.mem = MEM_HOST
        io.channels = ch
        io.left, io.left_step, io.left_frame_stride = L.ctypes.data, L.strides[1], L.strides[0]
        io.right, io.right_step, io.right_frame_stride = R.ctypes.data, R.strides[1], R.strides[0]
        disp = None
        if want_disp:
            disp = np.empty((B, self.H, self.W), np.int16)
            io.disp, io.disp_step, io.disp_frame_stride = disp.ctypes.data, disp.strides[1], disp.strides[0]
        st = stages or 15
        clouds = counts = None
        if st & (STAGE_CLOUD | STAGE_VOXEL | STAGE_SOR):
            clouds = np.empty((B, cap), POINT_DTYPE)
            counts = np.zeros(B, np.int32)
            io.cloud, io.cloud_capacity, io.cloud_count = clouds.ctypes.data, cap, counts.ctypes.data
        io.stages = st
        self._check(self.lib.cvr_process_batch(self.ctx, B, C.byref(io)))
        out = [clouds[i, :counts[i]].copy() for i in range(B)] if clouds is not None else None
        return disp, out

    def process_batch_raw(self, io, nframes):
        self._check(self.lib.cvr_process_batch(self.ctx, int(nframes), C.byref(io)))

    def submit_batch_raw(self, io, nframes):
        self._check(self.lib.cvr_submit_batch(self.ctx, int(nframes), C.byref(io)))

    def wait_batch(self):
        self._check(self.lib.cvr_wait_batch(self.ctx))

    # -- the gather of the filtered clouds to the mapping rank (MappingSystem::AddFrame -> MapDataBase::InsertBlock) --
    def comm_init(self, nranks, rank, unique_id):
        self._check(self.lib.cvr_comm_init(self.ctx, int(nranks), int(rank), C.c_char_p(bytes(unique_id))))
        self._nranks = int(nranks)

    def gather_clouds_raw(self, root, mem, out_ptr, out_capacity, frame_offsets_ptr, frame_offsets_capacity, frames_of_rank_ptr):
        self._check(self.lib.cvr_gather_clouds(self.ctx, int(root), int(mem), out_ptr, int(out_capacity), frame_offsets_ptr,
                                               int(frame_offsets_capacity), frames_of_rank_ptr))

    def GatherClouds(self, root=0, capacity=None, max_frames=None):
        """Gathers the clouds of the batch completed last to `root` (host memory).  Returns (points, frame_offsets,
        frames_of_rank) on the root, None elsewhere."""
        nr = getattr(self, "_nranks", 1)
        cap = capacity or self.max_points * self.params.max_batch
        mf = max_frames or nr * self.params.max_batch
        out = np.empty(cap, POINT_DTYPE)
        foff = np.zeros(mf + 1, np.int32)
        fr = np.zeros(nr, np.int32)
        self.gather_clouds_raw(root, MEM_HOST, _ptr(out), cap, _ptr(foff), len(foff), _ptr(fr))
        nf = int(fr.sum())
        return out[:foff[nf]].copy(), foff[:nf + 1].copy(), fr


def save_pcd(path, points):
    """pcl::io::savePCDFileBinary of a cloud (ReconstructionServer.cpp:165)."""
    pts = np.ascontiguousarray(points, POINT_DTYPE)
    rc = load_library().cvr_save_pcd(os.fsencode(path), _ptr(pts), len(pts))
    if rc != 0:
        raise CvrError(rc, load_library().cvr_last_error(None).decode())


def points_to_pcl(points):
    pts = np.ascontiguousarray(points, POINT_DTYPE)
    out = np.zeros((len(pts), 32), np.uint8)
    load_library().cvr_points_to_pcl(_ptr(pts), len(pts), _ptr(out))
    return out
