import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def cvr():
    """The product package (hyphenated directory name => importlib)."""
    return importlib.import_module("master-thesis_b200")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


def load_golden(name):
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))


GOLDEN_SMALL = ["s64x300_d64_bs5", "s64x300_d64_bs27", "s120x500_d64_bs27_hh", "s90x400_d48_bs9_minneg",
                "s90x400_d32_bs7_hh_minpos", "s100x420_d96_bs3"]
GOLDEN_LARGE = ["kitti_d128_bs27", "kitti_d128_bs9", "spi_d192_bs5_hh"]
# full-size BASELINE.json shapes at the reference's block size: CRC32 of cv2's outputs + every 16th disparity row
GOLDEN_BIG = ["hd_d256_bs27", "hd_d256_bs27_hh", "hd_d256_bs5", "spi_d192_bs27_hh"]
