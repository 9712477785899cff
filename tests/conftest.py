import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    """The .so files are git-ignored build products; (re)build what is missing. nvcc cross-compiles
    sm_100a without a GPU, so this also works in the CPU-only container."""
    import subprocess

    lib = os.path.join(ROOT, "another_cuda_sift_b200", "libb200sift.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "another_cuda_sift_b200", "csrc"), "-j8"],
                              stdout=subprocess.DEVNULL)
    cpu = os.path.join(ROOT, "oracle", "liboracle_sift.so")
    if not os.path.exists(cpu):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "cpu"], stdout=subprocess.DEVNULL)


@pytest.fixture(scope="session", autouse=True)
def built():
    _ensure_built()


def golden(name):
    import numpy as np

    path = os.path.join(GOLDEN, name)
    if not os.path.exists(path):
        pytest.skip(f"golden fixture {name} not generated yet (tests/golden/make_golden.py on a B200)")
    return np.load(path)
