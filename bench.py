#!/usr/bin/env python
"""bench.py -- stereo frames/sec through SGBM -> cloud -> filter (BASELINE.json metric) on N B200s.

A "step" = one pass of the hot path (cvtColor, StereoSGBM-exact disparity, prune, triangulate, voxel grid, statistical
outlier removal, then the gather of the filtered clouds to the mapping rank) over one batch of `--batch` KITTI-shape
synthetic pairs per GPU.  Workload = BASELINE.json configs[1]: KITTI-odometry-shape sequence, 1242x375, D=128,
config_default.json SGBM parameters, frames sharded over the ranks (frame-level data parallelism, weak scaling: per-GPU
batch fixed); the only exchange is cvr_gather_clouds (NCCL) to rank 0.

  value : frames/s with the batch already resident in HBM (device pointers through the C ABI), clouds gathered into a
          device buffer of rank 0
  e2e   : frames/s through the same C-ABI calls with pinned HOST buffers: H2D of both images inside the timed region, and
          the filtered clouds of EVERY rank delivered to rank 0's host memory (N = 1: the D2H copy of cvr_wait_batch)
  roofline     : dominant kernel (k_path_v4), algorithmic bytes / its CUDA-event time measured on the library's stream
                 during the timed steps; `path` = the whole step against SURVEY 8(d)'s 458 MB per frame;
                 `traffic` / `dram_bytes_per_frame` from the ncu capture committed under profiles/ (refused when the
                 capture was taken on other kernel sources or another batch size)
  side_shapes  : BASELINE.json configs[2], [3], [4] (1280x720 D=192 MODE_HH; 1920x1080 D=256; 20 M-point filter stress),
                 device-resident, each with its own clock sample and fraction of the HBM roofline (N = 1 only)
  cpu_baseline : the reference's CPU path on this box's host cores (cv2 = the OpenCV the reference calls; the PCL
                 half through the C port in oracle/), bounded sample, rank 0 at N=1 only
  --impl reference : that CPU path as its own arm.
"""
import argparse
import ctypes as C
import glob
import hashlib
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, D, BS = 1242, 375, 128, 27
WORKLOAD = "KITTI-odometry-shape sequence, synthetic rectified pairs 1242x375 BGR, StereoSGBM D=128 bs=27 (config_default.json), " \
           "prune + Triangulate3D + VoxelGrid(0.2) + SOR(k=30, 0.5), frames sharded over ranks"
UNIQUE_PAIRS = 16
# algorithmic bytes per frame (each stage boundary written once and read once)
ALG_BYTES = {"kitti": 458e6, "spi_hh": 2.14e9, "hd": 3.76e9, "filter_stress": 13.9e9}
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r02_traffic.json")


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def source_sha():
    """Hash of the kernel sources: an ncu capture only describes the kernels it was taken on."""
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "master-thesis_b200", "csrc", "*.cu*"))):
        h.update(os.path.basename(f).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.lines = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def make_pairs(n, h=H, w=W, d=D):
    synth = importlib.import_module("master-thesis_b200.synth")
    Ls, Rs = [], []
    for s in range(n):
        l, r = synth.synth_pair(h, w, d, seed=s, color=True)
        Ls.append(l); Rs.append(r)
    return np.stack(Ls), np.stack(Rs)


# ------------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own CPU implementation of the path
_PAIR_CACHE = {}


def _cpu_worker(args):
    seed, use_cv2 = args
    import numpy as _np
    synth = importlib.import_module("master-thesis_b200.synth")
    from oracle import oracle as O
    key = seed % UNIQUE_PAIRS
    if key not in _PAIR_CACHE:                       # input synthesis is not part of the timed path (cached after warm-up)
        _PAIR_CACHE[key] = synth.synth_pair(H, W, D, seed=key, color=True)
    lb, rb = _PAIR_CACHE[key]
    t0 = time.perf_counter()
    if use_cv2:
        import cv2
        cv2.setNumThreads(1)
        gl = cv2.cvtColor(lb, cv2.COLOR_BGR2GRAY); gr = cv2.cvtColor(rb, cv2.COLOR_BGR2GRAY)
        m = cv2.StereoSGBM_create(0, D, BS, P1=24 * BS * BS, P2=96 * BS * BS, disp12MaxDiff=0, preFilterCap=10, uniquenessRatio=10,
                                  speckleWindowSize=100, speckleRange=2, mode=cv2.STEREO_SGBM_MODE_SGBM)
        disp = m.compute(gl, gr)
        dn = cv2.normalize(disp, None, 0, 255, cv2.NORM_MINMAX, cv2.CV_8U)
        _, sd = cv2.meanStdDev(dn)
        _, mask = cv2.threshold(dn, sd[0][0] * 1.3, 255, cv2.THRESH_BINARY)
        pr = disp.copy(); pr[mask == 0] = 0
        dispf = pr.astype(_np.float32) * _np.float32(1 / 16)
        mask = mask.copy(); mask[dispf <= 0] = 0
    else:
        gl, gr = O.gray_bgr(lb), O.gray_bgr(rb)
        disp = O.sgbm(gl, gr, O.reference_sgbm_params(block_size=BS, num_disparities=D))
        pr, dispf, mask, _, _ = O.prune(disp)
    pts = O.triangulate(dispf, lb, mask, _np.float32(958.35805), _np.float32(0.532671), _np.float32(696.0217), _np.float32(224.1806))
    vox, _, _ = O.voxel_grid(pts, 0.2)
    kept, _, _, _ = O.sor(vox, 30, 0.5)
    return time.perf_counter() - t0, len(kept)


def cpu_reference_run(steps, warmup, frames_per_step=None):
    """Times the CPU path frame-parallel over all host cores (one frame per worker process; OpenCV's MODE_SGBM is
    single-threaded).  Returns (frames/s, cores, kind, sample description, ms_per_step)."""
    import multiprocessing as mp
    from oracle import oracle as O
    O.build()
    try:
        import cv2  # noqa: F401
        use_cv2 = True
    except Exception:
        use_cv2 = False
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    workers = max(1, min(cores, 256))
    fps_step = frames_per_step or workers
    ctx = mp.get_context("fork")
    with ctx.Pool(workers) as pool:
        for _ in range(warmup):
            pool.map(_cpu_worker, [(i, use_cv2) for i in range(workers)])
        t0 = time.perf_counter()
        n = 0
        for s in range(steps):
            res = pool.map(_cpu_worker, [(s * fps_step + i, use_cv2) for i in range(fps_step)])
            n += len(res)
        dt = time.perf_counter() - t0
    kind = "reference" if use_cv2 else "port"
    sample = "%d KITTI-shape frames (%d steps x %d), %d worker processes; %s" % (
        n, steps, fps_step, workers,
        "cv2 (the OpenCV the reference calls) for cvtColor+StereoSGBM+normalize/meanStdDev/threshold, C port (oracle/) for "
        "Triangulate3D + VoxelGrid + SOR since PCL is absent" if use_cv2 else "C port (oracle/) for the whole path")
    return n / dt, workers, kind, sample, dt / max(steps, 1) * 1e3


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    v, cores, kind, sample, ms = cpu_reference_run(args.steps, args.warmup)
    out = {
        "impl": "reference", "metric": "stereo frames/sec (SGBM->cloud->filter)", "value": v, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int16", "data": "synthetic", "config": {"workload": WORKLOAD},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------------------
# side shapes: BASELINE.json configs[2..4] on one GPU, device-resident
def _side_stereo(cvr, torch, dev, local, name, h, w, d, bs, mode, batch, steps, peak):
    cfg = cvr.parse_config(None)
    cfg.update(block_size=bs, num_disparities=d)
    rec = cvr.Reconstruct3D(w, h, config=cfg, mode=mode, max_batch=batch + batch // 4, device=local)
    B = rec.suggest_batch(batch)
    Ls, Rs = make_pairs(4, h, w, d)
    idx = [i % 4 for i in range(B)]
    dL = torch.from_numpy(Ls[idx]).to(dev); dR = torch.from_numpy(Rs[idx]).to(dev)
    cap = 200000 * B                                                   # packed points of a whole batch
    ios = []
    keep = []
    for _ in range(2):
        cl = torch.empty((cap, 16), dtype=torch.uint8, device=dev); off = torch.zeros(B + 1, dtype=torch.int32, device=dev)
        cnt = torch.zeros(B, dtype=torch.int32, device=dev)
        io = cvr.Batch()
        io.mem, io.channels, io.stages = cvr.MEM_DEVICE, 3, 15
        io.left, io.left_step, io.left_frame_stride = dL.data_ptr(), w * 3, w * h * 3
        io.right, io.right_step, io.right_frame_stride = dR.data_ptr(), w * 3, w * h * 3
        io.cloud_packed, io.cloud_packed_capacity, io.cloud_offsets, io.cloud_count = cl.data_ptr(), cap, off.data_ptr(), cnt.data_ptr()
        ios.append(io); keep.append((cl, off, cnt))

    def run(n):
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        rec.submit_batch_raw(ios[0], B)
        for i in range(1, n):
            rec.submit_batch_raw(ios[i & 1], B)
            rec.wait_batch()
        rec.wait_batch()
        e1.record()
        torch.cuda.synchronize()
        return max(e0.elapsed_time(e1), 0.9 * (time.perf_counter() - t0) * 1e3)
    run(3)
    clocks = ClockSampler(local)
    ms = run(steps)
    clk = clocks.stop()
    rec.close()
    fps = steps * B / (ms * 1e-3)
    key = "spi_hh" if name.startswith("spi") else "hd"
    alg = ALG_BYTES[key] if not (key == "hd" and mode == 1) else 6.5e9
    ach = alg * fps / 1e9
    return {"workload": "%dx%d D=%d bs=%d %s, prune + Triangulate3D + VoxelGrid(0.2) + SOR(30, 0.5), device-resident" % (w, h, d, bs, "MODE_HH" if mode else "MODE_SGBM"),
            "frames_per_s": fps, "batch": B, "steps": steps, "ms_per_step": ms / steps,
            "roofline": {"bound": "hbm", "alg_bytes_per_frame": alg, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak}, "clocks": clk}


def _side_filter_stress(cvr, torch, dev, local, n, peak):
    rng = np.random.default_rng(0)
    xyz = rng.random((n, 3), dtype=np.float32) * np.array([40, 12, 40], np.float32)
    no = n // 100
    # SURVEY 8(d) cfg5 puts the 1 % outliers in a 3x larger box, which overflows PCL's int32 voxel index at leaf 0.05 (the
    # filter would pass the cloud through); a 1.5x box keeps the grid (1200 x 360 x 1200) inside it
    xyz[:no] = (rng.random((no, 3), dtype=np.float32) - np.float32(1 / 6)) * np.array([60, 18, 60], np.float32)
    pts = np.zeros(n, cvr.POINT_DTYPE)
    pts["x"], pts["y"], pts["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    pts["r"] = rng.integers(0, 256, n); pts["g"] = 7; pts["b"] = 9; pts["a"] = 255
    cfg = cvr.parse_config(None)
    cfg.update(outlier_min_k=50, outlier_std_threshold=1.0)
    rec = cvr.Reconstruct3D(64, 64, config=cfg, voxel_leaf=0.05, max_points=n, filter_only=True, device=local)
    d_in = torch.from_numpy(pts.view(np.uint8).reshape(n, 16)).to(dev)
    d_vox = torch.empty((n, 16), dtype=torch.uint8, device=dev)
    d_out = torch.empty((n, 16), dtype=torch.uint8, device=dev)

    def run():
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m = rec.filter_cloud_device(d_in.data_ptr(), n, cvr.STAGE_VOXEL, d_vox.data_ptr(), n)
        t1 = time.perf_counter()
        tv = dict(rec.last_timings())
        k = rec.filter_cloud_device(d_vox.data_ptr(), m, cvr.STAGE_SOR, d_out.data_ptr(), n)
        t2 = time.perf_counter()
        ts = dict(rec.last_timings())
        return m, k, tv.get("voxel", 0.0), ts.get("sor", 0.0), (t1 - t0) * 1e3, (t2 - t1) * 1e3
    rec.set_profiling(True)
    run()
    clocks = ClockSampler(local)
    res = [run() for _ in range(3)]
    clk = clocks.stop()
    rec.close()
    m, k = res[0][0], res[0][1]
    vox_ms = float(np.mean([r[2] for r in res])); sor_ms = float(np.mean([r[3] for r in res]))
    call_ms = float(np.mean([r[4] + r[5] for r in res]))
    alg = 16.0 * n + 24.0 * n + 16.0 * m + m * (24.0 + 12 * 50) + 16.0 * k            # SURVEY 8(d) cfg5 formula with the measured M, K
    ach = alg / ((vox_ms + sor_ms) * 1e-3) / 1e9
    return {"workload": "%d-point cloud (uniform 40x12x40 m + 1 %% outliers), VoxelGrid leaf 0.05 then SOR k=50 alpha=1.0, device-resident" % n,
            "points_in": n, "voxels": m, "kept": k, "voxel_ms": vox_ms, "sor_ms": sor_ms, "both_calls_wall_ms": call_ms,
            "roofline": {"bound": "hbm", "alg_bytes": alg, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "floor_ms": alg / (peak * 1e9) * 1e3}, "clocks": clk}


def _side_png(cvr, torch, dev, local):
    """Ingest: the PNG files the reference's streamers send, decoded on the device (cvr_decode_png_batch,
    device output) next to cv2.imdecode on one host core.  A stream is decoded by one thread, so the rate of the device decoder
    is the number of files per call: measured at 220 (both cameras of one 110-frame batch) and at 2200 (ten batches)."""
    import cv2
    Ls, Rs = make_pairs(8)
    files = [cv2.imencode(".png", im)[1].tobytes() for im in list(Ls) + list(Rs)]      # OpenCV's default PNG settings, as the streamers use
    rates = {}
    ok = True
    clocks = ClockSampler(local)
    for n, reps in ((220, 5), (2200, 2)):
        # the received files sit in pinned host memory (a server's receive buffer), one after the other
        sizes = [len(files[i % len(files)]) for i in range(n)]
        pin = torch.empty(int(np.sum(sizes)), dtype=torch.uint8).pin_memory()
        arr = pin.numpy()
        batch, o = [], 0
        for i in range(n):
            arr[o:o + sizes[i]] = np.frombuffer(files[i % len(files)], np.uint8)
            batch.append(arr[o:o + sizes[i]])
            o += sizes[i]
        rec = cvr.Reconstruct3D(W, H, filter_only=True, max_points=16, device=local)
        out = torch.empty((n, H, W, 3), dtype=torch.uint8, device=dev)
        rec.DecodePNG(batch, device_out=out.data_ptr())
        ok = ok and bool(np.array_equal(out[n - 1].cpu().numpy(), cv2.imdecode(np.asarray(batch[n - 1]), cv2.IMREAD_COLOR)))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            rec.DecodePNG(batch, device_out=out.data_ptr())
        rates[n] = reps * n / (time.perf_counter() - t0)
        rec.close()
        del out, pin, arr, batch
        torch.cuda.empty_cache()
    clk = clocks.stop()
    cv2.setNumThreads(1)
    t0 = time.perf_counter()
    for f in files * 8:
        cv2.imdecode(np.frombuffer(f, np.uint8), cv2.IMREAD_COLOR)
    dth = time.perf_counter() - t0
    return {"workload": "PNG files (1242x375 BGR, cv2.imencode defaults, %.0f KB each) in pinned host memory -> BGR frames in device memory, "
                        "2200 files per call (220 per call: images_per_s_at_220)" % (np.mean([len(f) for f in files]) / 1e3),
            "images_per_s": rates[2200], "stereo_frames_per_s": rates[2200] / 2, "images_per_s_at_220": rates[220], "matches_cv2": ok,
            "host_cv2_imdecode_images_per_s_per_core": len(files) * 8 / dth, "clocks": clk}


# ------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=104, help="frames per GPU per step")
    ap.add_argument("--exact-batch", action="store_true", help="use --batch as given instead of letting the library round it to full cluster waves")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-side-shapes", action="store_true", help="skip BASELINE.json configs[2..4] (they are measured at N=1 only)")
    ap.add_argument("--stress-points", type=int, default=20_000_000)
    ap.add_argument("--rectify", action="store_true", help="side measurement: config requires_rectification = true (RectifyImages on the device before the matcher)")
    ap.add_argument("--matcher", default="sgbm", choices=["sgbm", "bm"], help="sgbm = the headline path; bm = the reference's other block matcher (side measurement)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    # stdout carries exactly one JSON line: native libraries that write to fd 1 (NCCL prints its version banner there)
    # are sent to stderr for the duration of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cvr = importlib.import_module("master-thesis_b200")
    B = args.batch
    cfg = cvr.parse_config(None)
    if args.matcher == "bm":
        cfg["block_matcher"] = cvr.MATCHER_BM
    cfg["requires_rectification"] = bool(args.rectify)
    # --batch is a target: the library rounds it to a multiple of the clusters its vertical path kernel runs at once on this
    # GPU; every rank asks its own GPU, all ranks then use the smallest answer
    rec = cvr.Reconstruct3D(W, H, config=cfg, max_batch=B + B // 4, device=local)   # defaults = config_default.json SGBM/SBM sections, leaf 0.2, SOR 30/0.5
    if args.rectify:                                                   # a KITTI-like raw calibration (5 distortion terms, small rotation)
        K = np.array([[959.791, 0, 696.0217], [0, 956.9251, 224.1806], [0, 0, 1]])
        Dk = [-0.3728755, 0.2037299, 0.002219027, 0.001383707, -0.07233722]
        a = 0.009
        Rr = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
        Pp = np.array([[880.0, 0, 660.0, 0], [0, 880.0, 200.0, 0], [0, 0, 1, 0]])
        rec.SetRectification(K, Dk, Rr, Pp, K, Dk, Rr, Pp)
    if not args.exact_batch:
        B = rec.suggest_batch(B)
        if world > 1:
            tb = torch.tensor([B], dtype=torch.int32, device=dev)
            dist.all_reduce(tb, op=dist.ReduceOp.MIN)
            B = int(tb.item())
    if world > 1:
        # the library's own communicator for cvr_gather_clouds: rank 0 makes the NCCL id, torch carries the 128 bytes
        idt = torch.zeros(cvr.COMM_ID_BYTES, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(cvr.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        rec.comm_init(world, rank, idt.cpu().numpy().tobytes())
    Ls, Rs = make_pairs(UNIQUE_PAIRS)
    idx = [(rank * B + i) % UNIQUE_PAIRS for i in range(B)]            # frame_id sharding; the pool of unique pairs is cycled
    hL = torch.from_numpy(Ls[idx]).pin_memory(); hR = torch.from_numpy(Rs[idx]).pin_memory()
    dL = hL.to(dev); dR = hR.to(dev)
    cap = 16384 * B                                                    # points per batch after voxel + SOR (~1.6k per frame observed)
    # two batches are in flight at a time: two sets of output buffers, used alternately
    d_clouds = [torch.empty((cap, 16), dtype=torch.uint8, device=dev) for _ in range(2)]
    d_offs = [torch.zeros(B + 1, dtype=torch.int32, device=dev) for _ in range(2)]
    h_clouds = [torch.empty((cap, 16), dtype=torch.uint8).pin_memory() for _ in range(2)]
    h_offs = [torch.zeros(B + 1, dtype=torch.int32).pin_memory() for _ in range(2)]
    h_cnts = [torch.zeros(B, dtype=torch.int32).pin_memory() for _ in range(2)]
    frame_bytes = W * H * 3
    gathered_bytes = [0]

    def batch_struct(mem, L, R, cloud, off, cnt):
        io = cvr.Batch()
        io.mem = mem; io.channels = 3
        io.left, io.left_step, io.left_frame_stride = L.data_ptr(), W * 3, frame_bytes
        io.right, io.right_step, io.right_frame_stride = R.data_ptr(), W * 3, frame_bytes
        if cloud is not None:
            io.cloud_packed, io.cloud_packed_capacity, io.cloud_offsets = cloud.data_ptr(), cap, off.data_ptr()
        if cnt is not None:
            io.cloud_count = cnt.data_ptr()
        io.stages = 15
        return io
    io_dev = [batch_struct(cvr.MEM_DEVICE, dL, dR, d_clouds[k], d_offs[k], None) for k in range(2)]
    if world == 1:
        io_host = [batch_struct(cvr.MEM_HOST, hL, hR, h_clouds[k], h_offs[k], h_cnts[k]) for k in range(2)]
    else:       # N > 1: no local download of the clouds -- they travel device to device to rank 0, which takes them to its host
        io_host = [batch_struct(cvr.MEM_HOST, hL, hR, None, None, h_cnts[k]) for k in range(2)]

    # the mapping rank's receive buffers for cvr_gather_clouds
    gcap = cap * world
    g_dev = torch.empty((gcap, 16), dtype=torch.uint8, device=dev) if (world > 1 and rank == 0) else None
    g_host = torch.empty((gcap, 16), dtype=torch.uint8).pin_memory() if (world > 1 and rank == 0) else None
    g_foff = np.zeros(B * world + 1, np.int32); g_fr = np.zeros(world, np.int32)

    def gather(to_host):
        if world == 1:
            return
        if rank == 0:
            buf = g_host if to_host else g_dev
            rec.gather_clouds_raw(0, cvr.MEM_HOST if to_host else cvr.MEM_DEVICE, buf.data_ptr(), gcap, g_foff.ctypes.data, len(g_foff), g_fr.ctypes.data)
            gathered_bytes[0] = int(g_foff[int(g_fr.sum())]) * 16
        else:
            rec.gather_clouds_raw(0, cvr.MEM_HOST if to_host else cvr.MEM_DEVICE, None, 0, None, 0, None)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(io, steps, profile=False, host=False):
        rec.set_profiling(profile)
        stage_ms = {}
        barrier()
        l0 = rec.launch_count()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        # Two batches in flight (cvr_submit_batch / cvr_wait_batch): the upload and the matcher of step i+1 overlap the
        # download of step i.  `host`: pinned host buffers (every step's H2D and D2H inside the timed region; the same
        # buffers are re-sent every step).  After each wait the step's clouds are gathered to rank 0 (cvr_gather_clouds,
        # on the library's communication stream, while the next batch computes).  With `profile` the library brackets
        # every stage with events on the stream it runs on and hands the durations back at the wait.
        def collected(i):
            gather(host)
            if profile:
                for n_, ms_ in rec.last_timings():
                    stage_ms.setdefault(n_, []).append(ms_)
        rec.submit_batch_raw(io[0], B)
        for i in range(1, steps):
            rec.submit_batch_raw(io[i & 1], B)
            rec.wait_batch()
            collected(i - 1)
        rec.wait_batch()
        collected(steps - 1)
        e1.record()
        torch.cuda.synchronize()
        # every library call blocks until its (own) stream is idle, so both torch-stream events are recorded with all
        # device work of the bracketed calls complete; wall time is kept as a cross-check
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        if ms < 0.9 * wall * 1e3:
            ms = wall * 1e3
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rec.set_profiling(False)
        return float(t.item()), rec.launch_count() - l0, stage_ms

    # warm-up (both paths), then the timed regions
    timed(io_dev, max(args.warmup, 3))
    timed(io_host, 2, host=True)
    clocks = ClockSampler(local) if rank == 0 else None
    ms_dev, launches, stage_ms = timed(io_dev, args.steps, profile=True)
    ms_e2e, _, _ = timed(io_host, args.steps, host=True)
    clk = clocks.stop() if clocks else None
    counts = h_cnts[(args.steps - 1) & 1].numpy().copy()
    frames = args.steps * B * world
    value = frames / (ms_dev * 1e-3)
    e2e_v = frames / (ms_e2e * 1e-3)
    vcfg = rec.path_kernel_config() if hasattr(rec, "path_kernel_config") else None
    rec.close()

    if rank == 0:
        peak, peak_kind = hbm_peak()
        cells = (W - D) * H * D
        alg_bytes = 6.0 * cells * B                 # k_path_v4 per launch: read C, read S, write S (int16 each)
        kv = stage_ms.get("path_v_down", [])
        if args.matcher == "bm":
            kv = []
        kv_ms = float(np.mean(kv)) if kv else None
        step_ms = ms_dev / args.steps
        roof = None
        if kv_ms:
            ach = alg_bytes / (kv_ms * 1e-3) / 1e9
            # measured DRAM traffic: the ncu capture committed under profiles/, valid only for these kernel sources and this batch
            traffic, dram_frame, note = None, None, None
            try:
                with open(TRAFFIC_FILE) as f:
                    tr = json.load(f)
                if tr.get("src_sha") != source_sha():
                    note = "profiles/r02_traffic.json was captured on other kernel sources: refused"
                elif int(tr.get("batch", -1)) != B:
                    note = "profiles/r02_traffic.json was captured at batch %s, this run uses %d: refused" % (tr.get("batch"), B)
                else:
                    kk = tr["kernels"].get("k_path_v4") or tr["kernels"]["k_path_v3"]
                    traffic = float(kk["dram_read"]) + float(kk["dram_write"])
                    dram_frame = float(tr["dram_bytes_per_frame"])
                    note = "dram__bytes_read + dram__bytes_write per launch, ncu capture profiles/r02/%s" % tr.get("source", "")
            except Exception as ex:      # noqa: BLE001
                note = "no usable capture: %r" % (ex,)
            path_ach = ALG_BYTES["kitti"] * B / (step_ms * 1e-3) / 1e9
            roof = {"bound": "hbm", "kernel": "k_path_v4 (fused down-right / down / down-left path aggregation, several columns per warp)", "achieved": ach, "peak": peak, "peak_kind": peak_kind,
                    "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_note": note, "alg_bytes_per_launch": alg_bytes, "ms_per_launch": kv_ms,
                    "share_of_step": kv_ms / step_ms,
                    "path": {"alg_bytes_per_frame": ALG_BYTES["kitti"], "achieved": path_ach, "frac": path_ach / peak,
                             "what": "SURVEY 8(d): every stage boundary written once and read once; whole step time"},
                    "dram_bytes_per_frame": dram_frame,
                    "dram_achieved": (dram_frame * B / (step_ms * 1e-3) / 1e9) if dram_frame else None}
        h2d = 2 * frame_bytes * B
        d2h = (int(counts.sum()) * 16 + 4 * (2 * B + 1)) if world == 1 else (gathered_bytes[0] + 4 * B)
        out = {
            "metric": "stereo frames/sec (SGBM->cloud->filter)", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int16", "data": "synthetic (%d unique seeded pairs cycled)" % UNIQUE_PAIRS,
            "config": {"workload": WORKLOAD + (" + RectifyImages (requires_rectification)" if args.rectify else ""), "frames_per_gpu_per_step": B, "l2": "inputs larger than L2: working set per step (cost volumes 214 MB/frame x batch) >> 126 MB L2",
                       "parallelism": "frames sharded, dp%d, cvr_gather_clouds (NCCL send/recv) of the filtered clouds to rank 0 in both arms" % world},
            "e2e": {"value": e2e_v, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps,
                    "delivery": "rank 0 host memory holds the clouds of all ranks" if world > 1 else "host memory (one packed copy per batch)"},
            "gpu_launches": int(launches),
            "launches_per_step": launches / args.steps,
            "clocks": clk,
            "roofline": roof,
            "stage_ms_per_step": {k: float(np.mean(v)) for k, v in stage_ms.items()},
            "stage_us_per_frame": {k: float(np.mean(v)) / B * 1e3 for k, v in stage_ms.items()},
            "points_per_frame_after_filter": float(counts.mean()),
        }
        if world == 1 and not args.no_side_shapes and args.matcher == "sgbm" and not args.rectify:
            side = {}
            for nm, fn in (("spi_hh_bs27", lambda: _side_stereo(cvr, torch, dev, local, "spi", 720, 1280, 192, 27, 1, 30, 4, peak)),
                           ("spi_hh_bs5", lambda: _side_stereo(cvr, torch, dev, local, "spi", 720, 1280, 192, 5, 1, 30, 4, peak)),
                           ("hd", lambda: _side_stereo(cvr, torch, dev, local, "hd", 1080, 1920, 256, 27, 0, 14, 4, peak)),
                           ("filter_stress", lambda: _side_filter_stress(cvr, torch, dev, local, args.stress_points, peak)),
                           ("png_ingest", lambda: _side_png(cvr, torch, dev, local))):
                try:
                    side[nm] = fn()
                except Exception as ex:      # noqa: BLE001
                    side[nm] = {"error": repr(ex)}
                torch.cuda.empty_cache()
            out["side_shapes"] = side
        if world == 1 and not args.no_cpu_baseline:
            # run in a fresh process: no fork after CUDA initialisation, no interference with the GPU timing above
            try:
                r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "2", "--warmup", "1"],
                                   capture_output=True, text=True, timeout=900)
                out["cpu_baseline"] = json.loads(r.stdout.strip().splitlines()[-1])["cpu_baseline"]
            except Exception as ex:      # noqa: BLE001
                out["cpu_baseline"] = {"value": None, "unit": "frames/s", "cores": 0, "kind": "port", "sample": "failed: %r" % (ex,)}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
