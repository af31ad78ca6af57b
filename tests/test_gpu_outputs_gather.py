"""GPU: the output side of the batched calls -- packed cloud outputs, capacity errors, the pipeline surviving a failed
batch -- and cvr_gather_clouds (one rank here; two ranks over NCCL in test_gather_two_ranks_nccl when the box has two
GPUs).  Reference hand-over being replaced: MappingSystem.cpp:344-345 -> MapDataBase::InsertBlock (MapDataBase.cpp:28-54)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
H, W, D, BS, B = 60, 330, 64, 9, 3


def _cfg():
    return dict(block_size=BS, num_disparities=D, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10,
                min_disparity=0, outlier_min_k=8, outlier_std_threshold=0.5)


def _batch(oracle, seed0):
    Ls, Rs = zip(*[oracle.synth_pair(H, W, D, seed=seed0 + i, color=True) for i in range(B)])
    return np.ascontiguousarray(np.stack(Ls)), np.ascontiguousarray(np.stack(Rs))


def _io(cvr, mem, L, R):
    io = cvr.Batch()
    io.mem, io.channels, io.stages = mem, 3, 15
    if mem == cvr.MEM_HOST:
        io.left, io.right = L.ctypes.data, R.ctypes.data
    else:
        io.left, io.right = L.data_ptr(), R.data_ptr()
    io.left_step = io.right_step = W * 3
    io.left_frame_stride = io.right_frame_stride = W * H * 3
    return io


def test_packed_host_output_equals_strided(cvr, oracle):
    rec = cvr.Reconstruct3D(W, H, config=_cfg(), max_batch=B + 1, voxel_leaf=0.05)
    L, R = _batch(oracle, 500)
    _, want = rec.ProcessFrames(L, R)
    total = sum(len(c) for c in want)
    assert total > 100
    packed = np.zeros(total + 7, cvr.POINT_DTYPE)
    off = np.full(B + 1, -1, np.int32)
    cnt = np.zeros(B, np.int32)
    io = _io(cvr, cvr.MEM_HOST, L, R)
    io.cloud_packed, io.cloud_packed_capacity, io.cloud_offsets, io.cloud_count = packed.ctypes.data, len(packed), off.ctypes.data, cnt.ctypes.data
    for call in ("sync", "pipe"):
        packed[:] = 0; off[:] = -1
        if call == "sync":
            rec.process_batch_raw(io, B)
        else:
            rec.submit_batch_raw(io, B); rec.wait_batch()
        assert off[0] == 0 and off[B] == total and list(cnt) == [len(c) for c in want]
        for i in range(B):
            assert np.array_equal(packed[off[i]:off[i + 1]], want[i]), (call, i)
    # a packed buffer one point short is refused, the offsets still tell the caller how much room is needed
    io.cloud_packed_capacity = total - 1
    off[:] = -1
    with pytest.raises(cvr.CvrError) as e:
        rec.process_batch_raw(io, B)
    assert e.value.code == -5 and off[B] == total
    rec.close()


def test_device_outputs_report_capacity_instead_of_truncating(cvr, oracle):
    import torch
    rec = cvr.Reconstruct3D(W, H, config=_cfg(), max_batch=B, voxel_leaf=0.05)
    L, R = _batch(oracle, 510)
    _, want = rec.ProcessFrames(L, R)
    big = max(len(c) for c in want)
    dev = torch.device("cuda", 0)
    dL, dR = torch.from_numpy(L).to(dev), torch.from_numpy(R).to(dev)
    for cap, ok in ((big, True), (big - 1, False)):
        clouds = torch.zeros((B, cap, 16), dtype=torch.uint8, device=dev)
        counts = torch.zeros(B, dtype=torch.int32, device=dev)
        io = _io(cvr, cvr.MEM_DEVICE, dL, dR)
        io.cloud, io.cloud_capacity, io.cloud_count = clouds.data_ptr(), cap, counts.data_ptr()
        for call in ("sync", "pipe"):
            def run():
                if call == "sync":
                    rec.process_batch_raw(io, B)
                else:
                    rec.submit_batch_raw(io, B); rec.wait_batch()
            if ok:
                run()
                c = counts.cpu().numpy()
                for i in range(B):
                    assert np.array_equal(clouds[i, :c[i]].cpu().numpy().reshape(-1).view(cvr.POINT_DTYPE), want[i])
            else:
                with pytest.raises(cvr.CvrError) as e:
                    run()
                assert e.value.code == -5
                assert list(counts.cpu().numpy()) == [len(c) for c in want]      # the true counts, so the caller can resize
    # packed device output
    total = sum(len(c) for c in want)
    packed = torch.zeros((total, 16), dtype=torch.uint8, device=dev)
    off = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    io = _io(cvr, cvr.MEM_DEVICE, dL, dR)
    io.cloud_packed, io.cloud_packed_capacity, io.cloud_offsets = packed.data_ptr(), total, off.data_ptr()
    rec.submit_batch_raw(io, B); rec.wait_batch()
    o = off.cpu().numpy()
    flat = packed.cpu().numpy().reshape(-1).view(cvr.POINT_DTYPE)
    for i in range(B):
        assert np.array_equal(flat[o[i]:o[i + 1]], want[i])
    rec.close()


def test_failed_batch_does_not_wedge_the_pipeline(cvr, oracle):
    rec = cvr.Reconstruct3D(W, H, config=_cfg(), max_batch=B, voxel_leaf=0.05)
    L, R = _batch(oracle, 520)
    _, want = rec.ProcessFrames(L, R)
    small = np.zeros((B, 4), cvr.POINT_DTYPE)
    cnt = np.zeros(B, np.int32)
    bad = _io(cvr, cvr.MEM_HOST, L, R)
    bad.cloud, bad.cloud_capacity, bad.cloud_count = small.ctypes.data, 4, cnt.ctypes.data
    cap = W * H
    clouds = np.zeros((B, cap), cvr.POINT_DTYPE)
    good = _io(cvr, cvr.MEM_HOST, L, R)
    good.cloud, good.cloud_capacity, good.cloud_count = clouds.ctypes.data, cap, cnt.ctypes.data
    for _ in range(3):                                            # more failures than there are slots
        rec.submit_batch_raw(bad, B)
        with pytest.raises(cvr.CvrError) as e:
            rec.wait_batch()
        assert e.value.code == -5
    rec.submit_batch_raw(good, B); rec.submit_batch_raw(good, B)
    rec.wait_batch(); rec.wait_batch()
    for i in range(B):
        assert cnt[i] == len(want[i]) and np.array_equal(clouds[i, :cnt[i]], want[i])
    rec.close()


def test_gather_single_rank_is_a_local_copy(cvr, oracle):
    rec = cvr.Reconstruct3D(W, H, config=_cfg(), max_batch=B + 2, voxel_leaf=0.05)
    with pytest.raises(cvr.CvrError):
        rec.GatherClouds()                                        # nothing completed yet
    L, R = _batch(oracle, 530)
    _, want = rec.ProcessFrames(L, R)
    pts, foff, fr = rec.GatherClouds(root=0, capacity=W * H * B)
    assert list(fr) == [B] and len(foff) == B + 1
    for i in range(B):
        assert np.array_equal(pts[foff[i]:foff[i + 1]], want[i])
    # after a pipelined batch the gather serves that batch
    L2, R2 = _batch(oracle, 540)
    _, want2 = rec.ProcessFrames(L2, R2)
    io = _io(cvr, cvr.MEM_HOST, L2, R2)
    cnt = np.zeros(B, np.int32)
    io.cloud_count = cnt.ctypes.data
    rec.submit_batch_raw(io, B); rec.wait_batch()
    pts, foff, fr = rec.GatherClouds(root=0, capacity=W * H * B)
    for i in range(B):
        assert np.array_equal(pts[foff[i]:foff[i + 1]], want2[i])
    # too small an output is an error, not a truncation
    out = np.zeros(3, cvr.POINT_DTYPE)
    foff2 = np.zeros(B + 1, np.int32); fr2 = np.zeros(1, np.int32)
    with pytest.raises(cvr.CvrError) as e:
        rec.gather_clouds_raw(0, cvr.MEM_HOST, out.ctypes.data, 3, foff2.ctypes.data, B + 1, fr2.ctypes.data)
    assert e.value.code == -5
    rec.close()


def test_gather_two_ranks_nccl():
    """Two processes, one GPU each, NCCL: rank 0 ends up with both ranks' clouds, frame boundaries intact."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run on a multi-GPU box: gpurun --gpus 2)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29517", os.path.join(ROOT, "tests", "gpu_gather_ranks.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "gather ok" in r.stdout


def test_tail_stream_option_gives_the_same_clouds(cvr, oracle, monkeypatch):
    """CVR_TAIL_STREAM=1: the stages after the path kernels run on a second stream beside the next batch's matcher.  Same
    results as the single-stream order, batch after batch, also when a synchronous call is made while batches are in flight."""
    L0, R0 = _batch(oracle, 600)
    L1, R1 = _batch(oracle, 610)
    ref = cvr.Reconstruct3D(W, H, config=_cfg(), max_batch=B, voxel_leaf=0.05)
    want = [ref.ProcessFrames(L0, R0)[1], ref.ProcessFrames(L1, R1)[1]]
    want_disp = ref.GenerateDisparityMap(L0[0], R0[0])
    ref.close()
    monkeypatch.setenv("CVR_TAIL_STREAM", "1")
    rec = cvr.Reconstruct3D(W, H, config=_cfg(), max_batch=B, voxel_leaf=0.05)
    monkeypatch.delenv("CVR_TAIL_STREAM")
    cap = W * H
    bufs = [(np.zeros((B, cap), cvr.POINT_DTYPE), np.zeros(B, np.int32)) for _ in range(2)]
    ios = []
    for k, (L, R) in enumerate(((L0, R0), (L1, R1))):
        io = _io(cvr, cvr.MEM_HOST, L, R)
        io.cloud, io.cloud_capacity, io.cloud_count = bufs[k][0].ctypes.data, cap, bufs[k][1].ctypes.data
        ios.append(io)
    rec.submit_batch_raw(ios[0], B)
    for i in range(1, 7):                                           # two batches in flight, alternating inputs
        rec.submit_batch_raw(ios[i & 1], B)
        rec.wait_batch()
        k = (i - 1) & 1
        for f in range(B):
            assert bufs[k][1][f] == len(want[k][f]) and np.array_equal(bufs[k][0][f, :bufs[k][1][f]], want[k][f]), (i, f)
        bufs[k][0][:] = 0; bufs[k][1][:] = 0
    # a synchronous entry point while a batch is still in flight: ordered after it, and the batch is unharmed
    d = rec.GenerateDisparityMap(L0[0], R0[0])
    assert np.array_equal(d, want_disp)
    rec.wait_batch()
    for f in range(B):
        assert bufs[0][1][f] == len(want[0][f]) and np.array_equal(bufs[0][0][f, :bufs[0][1][f]], want[0][f])
    rec.close()
