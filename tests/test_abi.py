"""The C-ABI boundary: libb200sift.so loads, exports every symbol include/b200sift.h declares, mirrors
the reference's config struct, and fails loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "b200sift.h")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"B200SIFT_API[^;(]*?\b(b200sift_\w+)\s*\(", src)))


def test_header_declares_expected_surface():
    syms = header_symbols()
    for must in ("b200sift_create", "b200sift_destroy", "b200sift_detect_host", "b200sift_detect_device_async",
                 "b200sift_copy_to_host", "b200sift_results", "b200sift_match_device", "b200sift_match_host",
                 "b200sift_last_error"):
        assert must in syms
    assert len(syms) >= 20


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(os.path.join(ROOT, "another_cuda_sift_b200", "libb200sift.so"))
    missing = [s for s in header_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_python_signatures_cover_header():
    from another_cuda_sift_b200 import _lib

    assert sorted(_lib.SIGNATURES) == header_symbols()
    _lib.load()  # resolves all of them


def test_no_torch_types_in_header():
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    assert "torch" not in src and "thrust" not in src and "std::" not in src


def test_default_config_matches_reference_struct():
    # sift_cuda/types/CudaSiftConfig.hh:3-14
    from another_cuda_sift_b200 import _lib

    lib = _lib.load()
    cfg = _lib.Config()
    lib.b200sift_default_config(C.byref(cfg))
    assert (cfg.col_width, cfg.row_width, cfg.numFeatures, cfg.numOctaveLayers) == (0, 0, 5000, 3)
    assert (cfg.contrastThreshould, cfg.edgeThreshould, cfg.sigma, cfg.upscale) == (0.04, 10.0, 1.6, 0)
    from another_cuda_sift_b200 import CudaSiftConfig

    py = CudaSiftConfig()
    assert (py.numFeatures, py.numOctaveLayers, py.contrastThreshould, py.edgeThreshould, py.sigma, py.upscale) == \
        (5000, 3, 0.04, 10.0, 1.6, False)


def test_argument_errors_without_touching_the_gpu():
    from another_cuda_sift_b200 import _lib

    lib = _lib.load()
    h = C.c_void_p()
    cfg = _lib.Config()
    lib.b200sift_default_config(C.byref(cfg))
    assert lib.b200sift_create(C.byref(cfg), C.byref(h)) == -1  # Detector.cu:22-25: size not set
    assert b"width or height" in lib.b200sift_last_error()
    cfg.col_width, cfg.row_width, cfg.upscale = 64, 64, 1
    assert lib.b200sift_create(C.byref(cfg), C.byref(h)) == -1  # CudaSiftConfig.hh:12
    assert lib.b200sift_match_launches(2000, 2000) >= 1


def test_unset_size_returns_false_like_reference():
    from another_cuda_sift_b200 import CudaSiftConfig, Detector

    assert Detector(CudaSiftConfig()).gpuWarmUpAndAllocate() is False


def test_fails_loudly_without_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import numpy as np

    from another_cuda_sift_b200 import B200SiftError, CudaSiftConfig, Detector, matchBruteForce

    det = Detector(CudaSiftConfig(col_width=64, row_width=64))
    with pytest.raises(B200SiftError, match="no CPU fallback"):
        det.gpuWarmUpAndAllocate()
    with pytest.raises(B200SiftError):
        matchBruteForce(np.zeros((4, 128), np.float16), 4, np.zeros((4, 128), np.float16), 4)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "another_cuda_sift_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hh", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("on-device oracle", ""), f


def test_library_load_sets_hardware_queue_default_without_overriding():
    """Loading libb200sift.so puts CUDA_DEVICE_MAX_CONNECTIONS=32 into the process environment (the engine's frames in
    flight need more than CUDA's 8 default work queues) and never replaces a value the user chose. Checked in fresh
    processes, through the C library itself (getenv from libc, not os.environ, which is a snapshot)."""
    import subprocess
    import sys

    from another_cuda_sift_b200 import _lib

    code = ("import ctypes, os, sys\n"
            "libc = ctypes.CDLL(None); libc.getenv.restype = ctypes.c_char_p\n"
            "ctypes.CDLL(sys.argv[1])\n"
            "print((libc.getenv(b'CUDA_DEVICE_MAX_CONNECTIONS') or b'').decode())\n")
    env = {k: v for k, v in os.environ.items() if k != "CUDA_DEVICE_MAX_CONNECTIONS"}
    out = subprocess.run([sys.executable, "-c", code, _lib.LIB_PATH], env=env, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == "32", (out.stdout, out.stderr[-300:])
    env["CUDA_DEVICE_MAX_CONNECTIONS"] = "12"
    out = subprocess.run([sys.executable, "-c", code, _lib.LIB_PATH], env=env, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == "12", (out.stdout, out.stderr[-300:])
