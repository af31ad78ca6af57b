"""CPU: the C-ABI library builds, loads and exports every symbol include/cvrecon_b200.h declares; no compute."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "cvrecon_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(cvr_[a-z0-9_]+)\s*\(", hdr)))


def test_header_symbols_all_exported(cvr):
    if not os.path.exists(cvr.LIB_PATH):
        cvr.build()
    lib = cvr.load_library()
    declared = _declared_symbols()
    assert len(declared) >= 15
    for s in declared:
        assert hasattr(lib, s), "libcvrecon_b200.so does not export %s" % s
    assert set(declared) == set(cvr.EXPORTS)
    assert b"sm_100a" in lib.cvr_version()


def test_struct_layout_matches_header(cvr):
    # cvr_point is 16 bytes, the SGBM block is 11 ints; ctypes mirrors must agree with the C compiler
    assert cvr.POINT_DTYPE.itemsize == 16
    assert C.sizeof(cvr.SgbmParams) == 44
    import subprocess, tempfile, textwrap
    src = textwrap.dedent("""
        #include <stdio.h>
        #include "cvrecon_b200.h"
        int main(void) { printf("%zu %zu %zu\\n", sizeof(cvr_params), sizeof(cvr_batch), sizeof(cvr_point)); return 0; }
    """)
    with tempfile.TemporaryDirectory() as td:
        cfile = os.path.join(td, "s.c")
        open(cfile, "w").write(src)
        exe = os.path.join(td, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), cfile, "-o", exe])
        a, b, c = [int(v) for v in subprocess.check_output([exe]).split()]
    assert a == C.sizeof(cvr.Params) and b == C.sizeof(cvr.Batch) and c == 16


def test_default_params_follow_reference_config(cvr):
    lib = cvr.load_library()
    p = cvr.Params()
    lib.cvr_default_params(C.byref(p), 1242, 375)
    s = p.sgbm
    # config_default.json:13-21 and Reconstruct3D.cpp:63-64
    assert (s.block_size, s.num_disparities, s.uniqueness_ratio, s.speckle_range, s.speckle_window_size, s.pre_filter_cap, s.min_disparity) == \
        (27, 128, 10, 2, 100, 10, 0)
    assert s.P1 == 8 * 3 * 27 * 27 and s.P2 == 32 * 3 * 27 * 27 and s.disp12_max_diff == 0 and s.mode == cvr.MODE_SGBM
    assert abs(p.Q[10] - 0.005 * (959.791 + 903.7596) / 2) < 1e-5 and p.Q[0] == 1 and p.Q[15] == 1
    assert p.sor_mean_k == 30 and p.sor_stddev_mul == 0.5 and abs(p.voxel_leaf[0] - 0.2) < 1e-7 and p.prune_sigma_factor == 1.3


def test_no_cpu_fallback(cvr):
    """Without a GPU the product must fail loudly, never compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(cvr.CvrError) as e:
        cvr.Reconstruct3D(320, 240)
    assert e.value.code in (-4, -3)


def test_invalid_params_rejected(cvr):
    lib = cvr.load_library()
    p = cvr.Params()
    lib.cvr_default_params(C.byref(p), 320, 240)
    ctx = C.c_void_p()
    p.sgbm.num_disparities = 100          # not a multiple of 16: OpenCV rejects it too
    assert lib.cvr_create(C.byref(p), 0, C.byref(ctx)) == -1
    p.sgbm.num_disparities = 512
    assert lib.cvr_create(C.byref(p), 0, C.byref(ctx)) == -2
    p.sgbm.num_disparities = 64
    p.sgbm.block_size = 4
    assert lib.cvr_create(C.byref(p), 0, C.byref(ctx)) == -1
    assert b"odd" in lib.cvr_last_error(None)


def test_product_never_touches_oracle():
    """The product path must not import, link or call anything under oracle/."""
    pkg = os.path.join(ROOT, "master-thesis_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "cvr_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f
