#!/usr/bin/env python
"""bench.py -- stereo frames/sec through SGBM -> cloud -> filter (BASELINE.json metric) on N B200s.

A "step" = one pass of the hot path (cvr_process_batch: cvtColor, StereoSGBM-exact disparity, prune, triangulate,
voxel grid, statistical outlier removal) over one batch of `--batch` KITTI-shape synthetic pairs per GPU.
Workload = BASELINE.json configs[1]: KITTI-odometry-shape sequence, 1242x375, D=128, config_default.json SGBM
parameters, frames sharded over the ranks (frame-level data parallelism, weak scaling: per-GPU batch fixed); the
only exchange is the gather of the filtered clouds to the mapping rank (rank 0) over NCCL.

  value : frames/s with the batch already resident in HBM (device pointers through the C ABI)
  e2e   : frames/s through the same C-ABI call with pinned HOST buffers (H2D of both images and D2H of the
          filtered clouds inside the timed region)
  roofline     : dominant kernel (k_path_v, the fused vertical/diagonal path aggregation), algorithmic bytes / its
                 CUDA-event time measured on the library's stream during the timed steps
  cpu_baseline : the reference's CPU path on this box's host cores (cv2 = the OpenCV the reference calls; the PCL
                 half through the C port in oracle/), bounded sample, rank 0 at N=1 only
  --impl reference : that CPU path as its own arm.
"""
import argparse
import ctypes as C
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


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


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


def make_pairs(n):
    synth = importlib.import_module("master-thesis_b200.synth")
    Ls, Rs = [], []
    for s in range(n):
        l, r = synth.synth_pair(H, W, D, seed=s, color=True)
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
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=104, help="frames per GPU per step")
    ap.add_argument("--exact-batch", action="store_true", help="use --batch as given instead of letting the library round it to full cluster waves")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    # GPU (26 on the B200s measured so far => 104); every rank asks its own GPU, all ranks then use the smallest answer
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
    Ls, Rs = make_pairs(UNIQUE_PAIRS)
    idx = [(rank * B + i) % UNIQUE_PAIRS for i in range(B)]            # frame_id sharding; the pool of unique pairs is cycled
    hL = torch.from_numpy(Ls[idx]).pin_memory(); hR = torch.from_numpy(Rs[idx]).pin_memory()
    dL = hL.to(dev); dR = hR.to(dev)
    cap = 65536                                                        # points per frame after voxel + SOR (~11k observed)
    # two batches are in flight at a time: two sets of output buffers, used alternately
    d_clouds = [torch.empty((B, cap, 16), dtype=torch.uint8, device=dev) for _ in range(2)]
    d_cnts = [torch.zeros(B, dtype=torch.int32, device=dev) for _ in range(2)]
    h_clouds = [torch.empty((B, cap, 16), dtype=torch.uint8).pin_memory() for _ in range(2)]
    h_cnts = [torch.zeros(B, dtype=torch.int32).pin_memory() for _ in range(2)]
    h_cnt = h_cnts[0]
    frame_bytes = W * H * 3

    def batch_struct(mem, L, R, cloud, cnt):
        io = cvr.Batch()
        io.mem = mem; io.channels = 3
        io.left, io.left_step, io.left_frame_stride = L.data_ptr(), W * 3, frame_bytes
        io.right, io.right_step, io.right_frame_stride = R.data_ptr(), W * 3, frame_bytes
        io.cloud, io.cloud_capacity, io.cloud_count = cloud.data_ptr(), cap, cnt.data_ptr()
        io.stages = 15
        return io
    io_dev = [batch_struct(cvr.MEM_DEVICE, dL, dR, d_clouds[k], d_cnts[k]) for k in range(2)]
    io_host = [batch_struct(cvr.MEM_HOST, hL, hR, h_clouds[k], h_cnts[k]) for k in range(2)]

    # gather of the filtered clouds to the mapping rank (rank 0): counts by all_gather, payload by grouped send/recv
    gather_buf = torch.empty((world, B * cap // 4, 16), dtype=torch.uint8, device=dev) if (world > 1 and rank == 0) else None

    def gather_clouds(k):
        if world == 1:
            return
        d_cnt, d_cloud = d_cnts[k], d_clouds[k]
        cnts = [torch.empty_like(d_cnt) for _ in range(world)]
        dist.all_gather(cnts, d_cnt)
        tot = [int(c.sum().item()) for c in cnts]
        # compact this rank's clouds (variable length per frame) into one contiguous payload
        mask = torch.arange(cap, device=dev)[None, :] < d_cnt[:, None]
        payload = d_cloud[mask]
        ops = []
        if rank == 0:
            for r in range(1, world):
                ops.append(dist.P2POp(dist.irecv, gather_buf[r, :tot[r]], r))
        else:
            ops.append(dist.P2POp(dist.isend, payload, 0))
        if ops:
            for w_ in dist.batch_isend_irecv(ops):
                w_.wait()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(io, steps, with_gather=True, profile=False, pipelined=False):
        rec.set_profiling(profile)
        stage_ms = {}
        barrier()
        l0 = rec.launch_count()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        # Two batches in flight (cvr_submit_batch / cvr_wait_batch): the upload and the matcher of step i+1 overlap the
        # filter stages of step i and the download of step i-1.  `pipelined` = host buffers (every step's H2D and D2H are
        # inside the timed region; the same pinned host buffers are re-sent every step; each rank delivers its filtered
        # clouds to its own host process).  Otherwise device-resident inputs and outputs, and the NCCL gather of each
        # step's clouds to the mapping rank follows its wait.  With `profile` the library brackets every stage with
        # events on the stream it runs on and hands the durations back at the wait.
        def collected(i):
            if with_gather and not pipelined:
                gather_clouds(i & 1)
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
    timed(io_dev, max(args.warmup, 3), profile=False)
    timed(io_host, 2, pipelined=True)
    clocks = ClockSampler(local) if rank == 0 else None
    ms_dev, launches, stage_ms = timed(io_dev, args.steps, profile=True)
    ms_e2e, _, _ = timed(io_host, args.steps, pipelined=True)
    clk = clocks.stop() if clocks else None
    counts = h_cnt.numpy().copy()
    frames = args.steps * B * world
    value = frames / (ms_dev * 1e-3)
    e2e_v = frames / (ms_e2e * 1e-3)

    if rank == 0:
        peak, peak_kind = hbm_peak()
        cells = (W - D) * H * D
        alg_bytes = 6.0 * cells * B                 # k_path_v3 per launch: read C, read S, write S (int16 each)
        kv = stage_ms.get("path_v_down", [])
        if args.matcher == "bm":
            kv = []
        kv_ms = float(np.mean(kv)) if kv else None
        roof = None
        if kv_ms:
            ach = alg_bytes / (kv_ms * 1e-3) / 1e9
            # DRAM traffic of this kernel from the committed ncu --set full capture (profiles/r01f_k_path_v3_raw.csv:
            # dram__bytes_read.sum + dram__bytes_write.sum = 22.245 + 11.087 GB for a 104-frame launch), scaled to this batch
            traffic = (22.245403e9 + 11.087477e9) / 104.0 * B
            roof = {"bound": "hbm", "kernel": "k_path_v3<4,false,false,false,true,1024>", "achieved": ach, "peak": peak, "peak_kind": peak_kind,
                    "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "alg_bytes_per_launch": alg_bytes, "ms_per_launch": kv_ms,
                    "share_of_step": kv_ms / (ms_dev / args.steps)}
        out = {
            "metric": "stereo frames/sec (SGBM->cloud->filter)", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int16", "data": "synthetic (%d unique seeded pairs cycled)" % UNIQUE_PAIRS,
            "config": {"workload": WORKLOAD + (" + RectifyImages (requires_rectification)" if args.rectify else ""), "frames_per_gpu_per_step": B, "l2": "inputs larger than L2: working set per step (cost volumes 214 MB/frame x batch) >> 126 MB L2",
                       "parallelism": "frames sharded, dp%d, NCCL gather of filtered clouds to rank 0" % world},
            "e2e": {"value": e2e_v, "unit": "frames/s", "h2d_bytes_per_step": 2 * frame_bytes * B, "d2h_bytes_per_step": int(counts.sum()) * 16 + 4 * B,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": roof,
            "stage_ms_per_step": {k: float(np.mean(v)) for k, v in stage_ms.items()},
            "points_per_frame_after_filter": float(counts.mean()),
        }
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
    rec.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
