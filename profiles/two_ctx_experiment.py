"""Experiment: do two contexts (two streams) overlap each other's kernels well enough to beat one context with the same
total batch?  usage: python profiles/two_ctx_experiment.py"""
import importlib, os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
cvr = importlib.import_module("master-thesis_b200"); synth = importlib.import_module("master-thesis_b200.synth")
W, H, D = 1242, 375, 128
pairs = [synth.synth_pair(H, W, D, seed=s, color=True) for s in range(8)]
def setup(B):
    rec = cvr.Reconstruct3D(W, H, max_batch=B)
    L = torch.from_numpy(np.stack([pairs[i % 8][0] for i in range(B)])).cuda(); R = torch.from_numpy(np.stack([pairs[i % 8][1] for i in range(B)])).cuda()
    cap = 65536
    cloud = torch.empty((B, cap, 16), dtype=torch.uint8, device="cuda"); cnt = torch.zeros(B, dtype=torch.int32, device="cuda")
    io = cvr.Batch(); io.mem = cvr.MEM_DEVICE; io.channels = 3
    io.left, io.left_step, io.left_frame_stride = L.data_ptr(), W * 3, W * H * 3
    io.right, io.right_step, io.right_frame_stride = R.data_ptr(), W * 3, W * H * 3
    io.cloud, io.cloud_capacity, io.cloud_count = cloud.data_ptr(), cap, cnt.data_ptr(); io.stages = 15
    return rec, io, (L, R, cloud, cnt)
def run(recs, B, steps):
    def work(rec, io):
        for _ in range(steps): rec.process_batch_raw(io, B)
    for rec, io, _ in recs: rec.process_batch_raw(io, B)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(rec, io)) for rec, io, _ in recs]
    [t.start() for t in th]; [t.join() for t in th]
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    return len(recs) * B * steps / dt
for (nctx, B) in ((1, 72), (2, 36), (2, 72), (3, 36), (4, 36)):
    recs = [setup(B) for _ in range(nctx)]
    print("contexts", nctx, "batch", B, "frames/s", round(run(recs, B, 6)), flush=True)
    for r in recs: r[0].close()
    del recs; torch.cuda.empty_cache()
