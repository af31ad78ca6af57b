"""CPU, world_size 2, gloo: the host-side logic of the multi-GPU path -- frame sharding by rank and the gatherv of
variable-length filtered clouds to the mapping rank (rank 0).  The placement arithmetic is the library's own
(cvr_gather_layout, the function cvr_gather_clouds uses between its ncclAllGather of the headers and its grouped
ncclSend/ncclRecv); gloo all_gather / send / recv stand in for the NCCL calls, CPU tensors for the device buffers."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, maxB, cap, q):
    sys.path.insert(0, ROOT)
    cvr = importlib.import_module("master-thesis_b200")
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    B = maxB - rank                                               # ranks may complete batches of different sizes
    cnt = rng.integers(0, cap, size=B).astype(np.int32)
    # this rank's packed payload, every point tagged with (rank, frame)
    payload = np.zeros((int(cnt.sum()), 16), np.uint8)
    o = 0
    for i in range(B):
        payload[o:o + cnt[i], 0] = rank
        payload[o:o + cnt[i], 1] = i
        payload[o:o + cnt[i], 2:] = rng.integers(0, 256, size=(int(cnt[i]), 14), dtype=np.uint8)
        o += int(cnt[i])
    hdr = np.zeros(2 + maxB, np.int32)                            # the record k_pack_clouds builds on the device
    hdr[0], hdr[1], hdr[2:2 + B] = B, cnt.sum(), cnt
    hdrs = [torch.empty(2 + maxB, dtype=torch.int32) for _ in range(world)]
    dist.all_gather(hdrs, torch.from_numpy(hdr))
    allh = torch.stack(hdrs).numpy()
    roff, foff, fr = cvr.gather_layout(allh)
    assert roff[rank + 1] - roff[rank] == len(payload)
    if rank == 0:
        out = np.zeros((int(roff[-1]), 16), np.uint8)
        out[roff[0]:roff[1]] = payload
        for r in range(1, world):
            t = torch.empty((int(roff[r + 1] - roff[r]), 16), dtype=torch.uint8)
            dist.recv(t, src=r)
            out[roff[r]:roff[r + 1]] = t.numpy()
        ok = int(fr.sum()) == sum(maxB - r for r in range(world)) and len(foff) == int(fr.sum()) + 1
        f = 0
        for r in range(world):
            for i in range(int(fr[r])):
                seg = out[foff[f]:foff[f + 1]]
                ok &= len(seg) == int(allh[r, 2 + i]) and bool((seg[:, 0] == r).all()) and bool((seg[:, 1] == i).all())
                f += 1
        q.put(("ok" if ok else "bad", [int(x) for x in roff]))
    else:
        dist.send(torch.from_numpy(payload), dst=0)
    dist.barrier()
    dist.destroy_process_group()


def test_frame_sharding_and_cloud_gatherv_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 5, 50, q)) for r in range(2)]
    for p in procs:
        p.start()
    res, roff = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == "ok" and len(roff) == 3


def test_gather_layout_rejects_corrupt_headers(cvr):
    h = np.zeros((2, 6), np.int32)
    h[0] = [2, 7, 3, 4, 0, 0]
    h[1] = [1, 5, 5, 0, 0, 0]
    roff, foff, fr = cvr.gather_layout(h)
    assert list(roff) == [0, 7, 12] and list(foff) == [0, 3, 7, 12] and list(fr) == [2, 1]
    h[1, 1] = 6                                                   # total disagrees with the counts
    with pytest.raises(cvr.CvrError):
        cvr.gather_layout(h)
    h[1] = [5, 5, 5, 0, 0, 0]                                     # more frames than the header can describe
    with pytest.raises(cvr.CvrError):
        cvr.gather_layout(h)


def test_sharding_covers_every_frame_once():
    # bench.py: rank r owns global frames r*B .. r*B+B-1 of each step
    B, world = 36, 8
    owned = sorted(r * B + i for r in range(world) for i in range(B))
    assert owned == list(range(B * world))
