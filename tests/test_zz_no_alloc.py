"""GPU: cvr_prune_noalloc -- the prune step for callers that may not allocate device memory (a fixed-latency frame
loop, a plug-in host that bans allocation after start-up, a process whose memory pool another library owns).

The new entry point works from a scratch array built into the library instead of a cvr_ctx workspace.  This test
calls it through the C ABI twice in a row, on different frames, and requires both results to equal what the
existing entry point (cvr_prune, which uses the context's allocated workspace) returns for the same input --
exactly, as everywhere else in this suite for integer/byte outputs and for the exact x/16 float conversion.
The second call proves that nothing left in the shared scratch by the first one leaks into the next result.
Collected last on purpose (file name), so a run that stops at the first failure has already run everything else;
it passes both on its own in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frames(oracle):
    """Two different disparity maps of one size (taken from the CPU oracle's SGBM on seeded pairs)."""
    H, W, D = 72, 360, 64
    out = []
    for seed, bs in ((3, 5), (9, 9)):
        l, r = oracle.synth_pair(H, W, D, seed)
        out.append(oracle.sgbm(l, r, oracle.reference_sgbm_params(block_size=bs, num_disparities=D)))
    assert not np.array_equal(out[0], out[1])
    return H, W, out


def test_prune_noalloc_matches_cvr_prune_twice_in_a_row(cvr, oracle):
    import torch
    lib = cvr.load_library()
    H, W, disps = _frames(oracle)
    cfg = dict(block_size=5, num_disparities=64, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10,
               min_disparity=0, outlier_min_k=30, outlier_std_threshold=0.5)
    rec = cvr.Reconstruct3D(W, H, config=cfg)                    # only for the existing entry point
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    d_pruned = torch.empty((H, W), dtype=torch.int16, device=dev)
    d_f32 = torch.empty((H, W), dtype=torch.float32, device=dev)
    d_mask = torch.empty((H, W), dtype=torch.uint8, device=dev)
    for disp in disps:                                           # two calls in a row sharing the built-in scratch
        want_pruned, want_f32, want_mask = rec.SetupFrame(disp)  # cvr_prune
        d_in = torch.from_numpy(disp).to(dev)
        d_pruned.fill_(-1); d_f32.fill_(-1); d_mask.fill_(7)
        torch.cuda.synchronize()
        rc = lib.cvr_prune_noalloc(d_in.data_ptr(), W, H, C.c_double(1.3), d_pruned.data_ptr(), d_f32.data_ptr(), d_mask.data_ptr(), None)
        assert rc == 0, lib.cvr_last_error(None)
        torch.cuda.synchronize()                                 # the call only enqueues; wait before reading
        assert np.array_equal(d_pruned.cpu().numpy(), want_pruned)
        assert np.array_equal(d_f32.cpu().numpy(), want_f32)
        assert np.array_equal(d_mask.cpu().numpy(), want_mask)
        assert want_mask.any() and not want_mask.all()           # the comparison is not vacuous
        # ... and both equal the CPU oracle's prune (itself pinned against cv2 by tests/test_oracle_golden.py)
        o_pruned, o_f32, o_mask, _, _ = oracle.prune(disp)
        assert np.array_equal(want_pruned, o_pruned) and np.array_equal(want_f32, o_f32) and np.array_equal(want_mask, o_mask)
    # optional outputs may be left out
    rc = lib.cvr_prune_noalloc(d_in.data_ptr(), W, H, C.c_double(1.3), None, None, d_mask.data_ptr(), None)
    torch.cuda.synchronize()
    assert rc == 0 and np.array_equal(d_mask.cpu().numpy(), want_mask)
    # in place (d_pruned == d_disp), the natural port of TrackingFrame::PruneDisparityImage
    rc = lib.cvr_prune_noalloc(d_in.data_ptr(), W, H, C.c_double(1.3), d_in.data_ptr(), None, None, None)
    torch.cuda.synchronize()
    assert rc == 0 and np.array_equal(d_in.cpu().numpy(), want_pruned)
    rec.close()


def test_prune_noalloc_matches_cv2_golden(cvr):
    """The allocation-free prune against the cv2 fixture of the KITTI-shape frame (pruned map and mask by value)."""
    import torch
    from conftest import load_golden
    lib = cvr.load_library()
    g = load_golden("kitti_d128_bs27")
    H, W = g["disp"].shape
    torch.cuda.set_device(0)
    d_in = torch.from_numpy(g["disp"]).cuda()
    d_pruned = torch.empty((H, W), dtype=torch.int16, device="cuda")
    d_mask = torch.empty((H, W), dtype=torch.uint8, device="cuda")
    for _ in range(2):
        d_pruned.fill_(-1); d_mask.fill_(7)
        rc = lib.cvr_prune_noalloc(d_in.data_ptr(), W, H, C.c_double(1.3), d_pruned.data_ptr(), None, d_mask.data_ptr(), None)
        assert rc == 0, lib.cvr_last_error(None)
        torch.cuda.synchronize()
        assert np.array_equal(d_pruned.cpu().numpy(), g["pruned"]) and np.array_equal(d_mask.cpu().numpy(), g["mask"])


def test_prune_noalloc_refuses_frames_beyond_its_scratch(cvr):
    import torch
    lib = cvr.load_library()
    torch.cuda.set_device(0)
    d = torch.zeros(16, dtype=torch.int16, device="cuda")
    # 4096 x 1025 pixels is one row more than the 4 Mi-pixel scratch can hold: refused before anything is launched
    rc = lib.cvr_prune_noalloc(d.data_ptr(), 4096, 1025, C.c_double(1.3), None, None, None, None)
    assert rc == -5
    assert b"CVR_NOALLOC_MAX_PIXELS" in lib.cvr_last_error(None)
    assert lib.cvr_prune_noalloc(None, 16, 1, C.c_double(1.3), None, None, None, None) == -1
