"""CPU, world_size 2, gloo: the host-side logic of the multi-GPU path -- frame sharding by rank and the gatherv of
variable-length filtered clouds to the mapping rank (rank 0) -- exactly the torch.distributed calls bench.py issues
over NCCL, with CPU tensors standing in for the device buffers."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, B, cap, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    # this rank's frames: global frame ids rank*B .. rank*B+B-1 (bench.py's sharding), variable cloud sizes
    cnt = torch.from_numpy(rng.integers(0, cap, size=B).astype(np.int32))
    cloud = torch.zeros((B, cap, 16), dtype=torch.uint8)
    for i in range(B):
        cloud[i, :cnt[i]] = torch.from_numpy(rng.integers(0, 256, size=(int(cnt[i]), 16), dtype=np.uint8))
        cloud[i, :cnt[i], 0] = rank * B + i                      # tag every point with its global frame id
    cnts = [torch.empty_like(cnt) for _ in range(world)]
    dist.all_gather(cnts, cnt)
    tot = [int(c.sum()) for c in cnts]
    mask = torch.arange(cap)[None, :] < cnt[:, None]
    payload = cloud[mask]
    assert payload.shape[0] == tot[rank]
    if rank == 0:
        bufs = {0: payload}
        ops = []
        for r in range(1, world):
            bufs[r] = torch.empty((tot[r], 16), dtype=torch.uint8)
            ops.append(dist.P2POp(dist.irecv, bufs[r], r))
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        ok = True
        for r in range(world):
            off = 0
            for i in range(B):
                n = int(cnts[r][i])
                seg = bufs[r][off:off + n]
                ok &= bool((seg[:, 0] == r * B + i).all())       # frame order and boundaries survive the gather
                off += n
            ok &= off == tot[r]
        q.put(("ok" if ok else "bad", tot))
    else:
        for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, payload, 0)]):
            w.wait()
    dist.barrier()
    dist.destroy_process_group()


def test_frame_sharding_and_cloud_gatherv_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 4, 50, q)) for r in range(2)]
    for p in procs:
        p.start()
    res, tot = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == "ok" and len(tot) == 2


def test_sharding_covers_every_frame_once():
    # bench.py: rank r owns global frames r*B .. r*B+B-1 of each step
    B, world = 36, 8
    owned = sorted(r * B + i for r in range(world) for i in range(B))
    assert owned == list(range(B * world))
