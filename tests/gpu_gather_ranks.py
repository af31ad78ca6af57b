"""Launched by tests/test_gpu_outputs_gather.py::test_gather_two_ranks_nccl (and by hand:
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/gpu_gather_ranks.py)
One process per GPU.  Every rank runs the path on its own frames (rank r completes a batch of B - r frames), then
cvr_gather_clouds brings all clouds to rank 0 over NCCL; rank 0 recomputes the other ranks' frames locally and checks
every gathered frame, to host memory and to device memory, after a synchronous and after a pipelined batch."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
H, W, D, BS, B = 60, 330, 64, 9, 4


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cvr = importlib.import_module("master-thesis_b200")
    from oracle import oracle as O
    O.build()
    cfg = dict(block_size=BS, num_disparities=D, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10,
               min_disparity=0, outlier_min_k=8, outlier_std_threshold=0.5)
    rec = cvr.Reconstruct3D(W, H, config=cfg, max_batch=B, voxel_leaf=0.05, device=local)
    # the 128-byte NCCL id travels from rank 0 to the others by the host's own means (here: a torch broadcast)
    idt = torch.zeros(cvr.COMM_ID_BYTES, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(cvr.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    rec.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()))

    def frames_of(r, rnd):
        n = max(B - r, 1)
        Ls, Rs = zip(*[O.synth_pair(H, W, D, seed=1000 * rnd + 10 * r + i, color=True) for i in range(n)])
        return np.ascontiguousarray(np.stack(Ls)), np.ascontiguousarray(np.stack(Rs))

    ok = True
    for rnd, mode in enumerate(("sync_host", "pipe_device")):
        want = None
        if rank == 0:                                            # what every rank is going to produce
            want = []
            for r in range(world):
                L, R = frames_of(r, rnd)
                want += rec.ProcessFrames(L, R)[1]
        L, R = frames_of(rank, rnd)
        n = L.shape[0]
        if mode == "sync_host":
            rec.ProcessFrames(L, R)
            res = rec.GatherClouds(root=0, capacity=W * H * B * world)
            if rank == 0:
                pts, foff, fr = res
        else:
            io = cvr.Batch()
            io.mem, io.channels, io.stages = cvr.MEM_HOST, 3, 15
            io.left, io.right = L.ctypes.data, R.ctypes.data
            io.left_step = io.right_step = W * 3
            io.left_frame_stride = io.right_frame_stride = W * H * 3
            cnt = np.zeros(n, np.int32)
            io.cloud_count = cnt.ctypes.data
            rec.submit_batch_raw(io, n); rec.wait_batch()
            cap = W * H * B * world
            out = torch.zeros((cap, 16), dtype=torch.uint8, device=dev)
            foff = np.zeros(B * world + 1, np.int32); fr = np.zeros(world, np.int32)
            rec.gather_clouds_raw(0, cvr.MEM_DEVICE, out.data_ptr() if rank == 0 else None, cap if rank == 0 else 0,
                                  foff.ctypes.data if rank == 0 else None, len(foff), fr.ctypes.data if rank == 0 else None)
            if rank == 0:
                pts = out.cpu().numpy().reshape(-1).view(cvr.POINT_DTYPE)
        if rank == 0:
            ok &= list(fr) == [max(B - r, 1) for r in range(world)]
            for i, w in enumerate(want):
                ok &= bool(np.array_equal(pts[foff[i]:foff[i + 1]], w))
            ok &= sum(len(w) for w in want) == int(foff[len(want)]) > 0
    t = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    rec.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("gather ok" if int(t.item()) == 1 else "gather MISMATCH", flush=True)
    return 0 if int(t.item()) == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
