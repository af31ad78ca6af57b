"""CPU: the sequential inflate routine of the device PNG decoder (master-thesis_b200/csrc/k_png.cu, a
__host__ __device__ function: inflate_all, the code lane 0 of the kernel runs) compiled for the host
(tests/png_host_harness.cu), against zlib on streams that use every block type, long matches and long Huffman codes."""
import ctypes as C
import os
import shutil
import subprocess
import zlib

import numpy as np
import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def harness():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    out = os.path.join(ROOT, "tests", "_build", "libpng_host.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call([nvcc, "-O1", "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-w", "-o", out, os.path.join(ROOT, "tests", "png_host_harness.cu")])
    return C.CDLL(out)


def _inflate(L, z, n):
    out = np.zeros(n + 8, np.uint8)
    rc = L.png_inflate_host(z, C.c_longlong(len(z)), out.ctypes.data_as(C.c_void_p), C.c_longlong(n))
    return rc, bytes(out[:n])


@pytest.mark.parametrize("strategy", [zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FILTERED])
@pytest.mark.parametrize("level", [0, 1, 6, 9])
def test_inflate_matches_zlib(harness, level, strategy):
    rng = np.random.default_rng(level * 10 + strategy)
    noise = rng.integers(0, 256, 150000, dtype=np.uint8).tobytes()
    smooth = (np.cumsum(rng.integers(-2, 3, 250000)) % 256).astype(np.uint8).tobytes()
    rep = (b"abcdefgh" * 5000 + bytes(40000) + noise[:30000]) * 3          # matches at distances up to the full 32 KiB window
    skew = rng.choice(256, 200000, p=np.r_[[0.9], np.full(255, 0.1 / 255)]).astype(np.uint8).tobytes()   # long Huffman codes
    for raw in (noise, smooth, rep, skew, b"", b"x"):
        co = zlib.compressobj(level, zlib.DEFLATED, 15, 9, strategy)
        z = co.compress(raw) + co.flush()
        rc, got = _inflate(harness, z, len(raw))
        assert rc == 0 and got == raw


def test_inflate_rejects_damage(harness):
    raw = bytes(range(256)) * 400
    z = bytearray(zlib.compress(raw, 6))
    assert _inflate(harness, bytes(z), len(raw))[0] == 0
    assert _inflate(harness, bytes(z), len(raw) - 5)[0] == 5             # more output than the image has room for
    assert _inflate(harness, bytes(z[: len(z) // 2]), len(raw))[0] != 0   # truncated
    z[0] = 0x79
    assert _inflate(harness, bytes(z), len(raw))[0] == 1                  # not a zlib header
    rng = np.random.default_rng(1)
    for _ in range(200):                                                  # random damage never hangs or writes out of bounds
        d = bytearray(zlib.compress(raw, 6))
        k = int(rng.integers(2, len(d)))
        d[k] ^= 1 << int(rng.integers(0, 8))
        _inflate(harness, bytes(d), len(raw))
