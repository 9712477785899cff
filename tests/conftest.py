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


def golden_frames(g):
    """The input frames of a golden capture: stored uint8 copies for small frames, regenerated from the
    seed (and checked against the stored SHA-1) for big ones."""
    import hashlib

    import numpy as np

    from another_cuda_sift_b200.synth import next_frame, synth_frame

    rows, cols, n = int(g["rows"]), int(g["cols"]), int(g["nframes"])
    regen = [synth_frame(rows, cols, 0, density=float(g["density"]))]
    for fr in range(1, n):
        regen.append(next_frame(regen[-1], fr))
    out = []
    for fr in range(n):
        if f"f{fr}_img_u8" in g:
            stored = g[f"f{fr}_img_u8"].astype(np.float32)
            assert np.array_equal(stored, regen[fr]), "synthetic generator drifted from the golden capture"
            out.append(stored)
        else:
            assert hashlib.sha1(regen[fr].tobytes()).hexdigest() == str(g[f"f{fr}_img_sha1"]), \
                "synthetic generator drifted from the golden capture"
            out.append(regen[fr])
    return out
