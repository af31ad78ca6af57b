"""CPU: the C++ host mirror (master-thesis_b200/host/Reconstruct3D_b200.hpp) compiles against the C ABI with g++ and
fails loudly without a GPU; config parsing mirrors the reference's JSON keys.  GPU: the C++ example runs one frame
and agrees with the Python harness."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT


def _build_example(tmp_path, cvr, name="example_frame"):
    if not os.path.exists(cvr.LIB_PATH):
        cvr.build()
    exe = str(tmp_path / name)
    libdir = os.path.dirname(cvr.LIB_PATH)
    subprocess.check_call(["g++", "-std=c++14", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "master-thesis_b200", "host", name + ".cpp"), "-L", libdir, "-lcvrecon_b200",
                           "-Wl,-rpath," + libdir, "-o", exe])
    return exe


def test_cpp_gather_example_compiles(tmp_path, cvr):
    _build_example(tmp_path, cvr, "example_gather")


def test_cpp_mirror_compiles_and_has_no_cpu_fallback(tmp_path, cvr):
    exe = _build_example(tmp_path, cvr)
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, "320", "240", "/dev/zero", "/dev/zero", str(tmp_path / "d"), str(tmp_path / "c")], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr


def test_config_keys_follow_reference_json(cvr):
    ref_json = {"config": {"server": {"port": 7000}, "reconstruction": {
        "requires_rectification": False, "block_matcher": "stereo_sgbm", "SBM": {"window_size": 27, "num_disparities": 128},
        "SGBM": {"block_size": 27, "num_disparities": 128, "uniqueness_ratio": 10, "speckle_range": 2, "speckle_window_size": 100,
                 "pre_filter_cap": 10, "min_disparity": 0.0}},
        "point_cloud_post_processing": {"outlier_min_k": 30, "outlier_std_threshold": 0.5, "keypoint_detector": "SIFT", "feature_detector": "FPFH"}}}
    c = cvr.parse_config(ref_json)
    assert c == dict(requires_rectification=False, block_matcher=cvr.MATCHER_SGBM, window_size=27, bm_num_disparities=128, block_size=27, num_disparities=128,
                     uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10, min_disparity=0, outlier_min_k=30,
                     outlier_std_threshold=0.5)
    ref_json["config"]["reconstruction"]["block_matcher"] = "stereo_bm"      # the reference's default
    assert cvr.parse_config(ref_json)["block_matcher"] == cvr.MATCHER_BM
    ref_json["config"]["reconstruction"]["block_matcher"] = "something_else"  # ConfigParser.cpp:48-50 falls back to the block matcher
    assert cvr.parse_config(ref_json)["block_matcher"] == cvr.MATCHER_BM


def test_points_to_pcl_layout(cvr):
    pts = np.zeros(3, cvr.POINT_DTYPE)
    pts["x"], pts["y"], pts["z"] = [1, 2, 3], [4, 5, 6], [7, 8, 9]
    pts["b"], pts["g"], pts["r"], pts["a"] = 10, 20, 30, 255
    raw = cvr.points_to_pcl(pts)
    assert raw.shape == (3, 32)
    f = raw[:, :16].copy().view(np.float32).reshape(3, 4)
    assert np.array_equal(f[:, 0], [1, 2, 3]) and (f[:, 3] == 1.0).all()          # pcl::PointXYZRGB: x,y,z,1.0f
    assert (raw[:, 16] == 10).all() and (raw[:, 17] == 20).all() and (raw[:, 18] == 30).all() and (raw[:, 19] == 255).all()
    assert (raw[:, 20:] == 0).all()


def test_save_pcd_is_a_binary_pcd_of_pcl_pointxyzrgb(tmp_path, cvr):
    """cvr_save_pcd: the file pcl::io::savePCDFileBinary writes for a PointXYZRGB cloud -- header fields, then 16 bytes per point."""
    rng = np.random.default_rng(3)
    pts = np.zeros(257, cvr.POINT_DTYPE)
    for k in "xyz":
        pts[k] = rng.standard_normal(257).astype(np.float32)
    for k in "bgr":
        pts[k] = rng.integers(0, 256, 257)
    pts["a"] = 255
    path = tmp_path / "cloud.pcd"
    cvr.save_pcd(str(path), pts)
    raw = path.read_bytes()
    head, _, body = raw.partition(b"DATA binary\n")
    fields = dict(l.split(" ", 1) for l in head.decode().splitlines() if l and not l.startswith("#"))
    assert fields["VERSION"] in ("0.7", ".7") and fields["FIELDS"] == "x y z rgb" and fields["SIZE"] == "4 4 4 4"
    assert fields["TYPE"] == "F F F F" and fields["COUNT"] == "1 1 1 1" and fields["HEIGHT"] == "1"
    assert int(fields["WIDTH"]) == int(fields["POINTS"]) == 257 and len(body) == 257 * 16
    back = np.frombuffer(body, cvr.POINT_DTYPE)
    assert np.array_equal(back, pts)
    rgb = np.frombuffer(body, np.uint32).reshape(-1, 4)[:, 3]      # PCL's packed colour word: 0xAARRGGBB
    assert np.array_equal(rgb, (255 << 24) | (pts["r"].astype(np.uint32) << 16) | (pts["g"].astype(np.uint32) << 8) | pts["b"])
    cvr.save_pcd(str(tmp_path / "empty.pcd"), pts[:0])
    assert b"POINTS 0" in (tmp_path / "empty.pcd").read_bytes()
    with pytest.raises(cvr.CvrError):
        cvr.save_pcd(str(tmp_path / "no_such_dir" / "x.pcd"), pts)


@pytest.mark.gpu
def test_cpp_example_matches_python_harness(tmp_path, cvr, oracle):
    exe = _build_example(tmp_path, cvr)
    H, W, D, bs = 80, 360, 64, 9
    lb, rb = oracle.synth_pair(H, W, D, seed=21, color=True)
    lb.tofile(tmp_path / "l.bgr"); rb.tofile(tmp_path / "r.bgr")
    r = subprocess.run([exe, str(W), str(H), str(tmp_path / "l.bgr"), str(tmp_path / "r.bgr"), str(tmp_path / "d.i16"), str(tmp_path / "c.bin"),
                        str(bs), str(D), str(tmp_path / "bm.i16")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    disp = np.fromfile(tmp_path / "d.i16", np.int16).reshape(H, W)
    ref = oracle.sgbm(oracle.gray_bgr(lb), oracle.gray_bgr(rb), oracle.reference_sgbm_params(block_size=bs, num_disparities=D))
    assert np.array_equal(disp, ref)
    cloud = np.fromfile(tmp_path / "c.bin", cvr.POINT_DTYPE)
    assert int(r.stdout.split()[1]) == len(cloud)
    # SetBlockMatcherType(STEREO_BLOCK_MATCHER) + SetStereoBMNumDisparities + SetStereoBMWindowSize (Reconstruct3D.cpp:364-388)
    bm = np.fromfile(tmp_path / "bm.i16", np.int16).reshape(H, W)
    assert np.array_equal(bm, oracle.stereobm(oracle.gray_bgr(lb), oracle.gray_bgr(rb), oracle.reference_bm_params(bs, D)))
    uv = [float(v) for v in r.stdout.split("uv")[1].split()]
    f = np.float32((np.float32(959.791) + np.float32(956.9251)) / np.float32(2)); z = f * np.float32(0.532671) / np.float32(16)
    assert np.allclose(uv, [(np.float32(100) - np.float32(696.0217)) * (z / f), -(np.float32(50) - np.float32(224.1806)) * (z / f), z], rtol=1e-5)


def _gather_frames(tmp_path, oracle, world, n, H, W, D):
    with open(tmp_path / "frames.bgr", "wb") as f:
        for p in range(world * n):
            lb, rb = oracle.synth_pair(H, W, D, seed=900 + p, color=True)
            f.write(lb.tobytes()); f.write(rb.tobytes())
    return str(tmp_path / "frames.bgr")


@pytest.mark.gpu
def test_cpp_gather_example_one_rank(tmp_path, cvr, oracle):
    """A C++ host runs a batch through ProcessStereoFrames and gathers its map blocks (one rank: a local copy)."""
    exe = _build_example(tmp_path, cvr, "example_gather")
    H, W, D, n = 60, 330, 64, 3
    frames = _gather_frames(tmp_path, oracle, 1, n, H, W, D)
    r = subprocess.run([exe, str(W), str(H), str(n), frames], capture_output=True, text=True)
    assert r.returncode == 0 and "gather ok ranks 1 frames 3" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_cpp_gather_example_two_ranks(tmp_path, cvr, oracle):
    """Two C++ processes, one GPU each, exchanging the NCCL id through a file; rank 0 checks every gathered frame."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    exe = _build_example(tmp_path, cvr, "example_gather")
    H, W, D, n = 60, 330, 64, 3
    frames = _gather_frames(tmp_path, oracle, 2, n, H, W, D)
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), CVR_ID_FILE=str(tmp_path / "nccl.id"))
        procs.append(subprocess.Popen([exe, str(W), str(H), str(n), frames], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=300) for p in procs]
    assert procs[0].returncode == 0 and procs[1].returncode == 0, outs
    assert "gather ok ranks 2 frames 6" in outs[0][0]
