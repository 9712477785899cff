"""ctypes loader for libb200sift.so (CUDA kernels + C ABI, include/b200sift.h).

The library is built in-tree by ``another_cuda_sift_b200/csrc/Makefile`` (nvcc, sm_100a). There
is no fallback of any kind: if the shared object is missing, or a call fails, an exception is
raised -- nothing here ever routes to a CPU implementation.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200sift.so")
CSRC = os.path.join(HERE, "csrc")


class B200SiftError(RuntimeError):
    pass


class Config(C.Structure):
    """b200sift_config == CudaSiftConfig (reference sift_cuda/types/CudaSiftConfig.hh:3-14)."""

    _fields_ = [
        ("col_width", C.c_int),
        ("row_width", C.c_int),
        ("numFeatures", C.c_int),
        ("numOctaveLayers", C.c_int),
        ("contrastThreshould", C.c_double),
        ("edgeThreshould", C.c_double),
        ("sigma", C.c_double),
        ("upscale", C.c_int),
    ]


class Options(C.Structure):
    """b200sift_options (include/b200sift.h): engine options without a counterpart in CudaSiftConfig."""

    _fields_ = [
        ("frames_in_flight", C.c_int),
        ("remove_duplicates", C.c_int),
        ("interpolate_orientation", C.c_int),
        ("working_upscale", C.c_int),
        ("reserved", C.c_int * 4),
    ]


# name -> (restype, argtypes); every symbol include/b200sift.h declares
SIGNATURES = {
    "b200sift_default_config": (None, [C.POINTER(Config)]),
    "b200sift_last_error": (C.c_char_p, []),
    "b200sift_version": (C.c_char_p, []),
    "b200sift_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "b200sift_default_options": (None, [C.POINTER(Options)]),
    "b200sift_create_ex": (C.c_int, [C.POINTER(Config), C.POINTER(Options), C.POINTER(C.c_void_p)]),
    "b200sift_destroy": (C.c_int, [C.c_void_p]),
    "b200sift_num_octaves": (C.c_int, [C.c_void_p]),
    "b200sift_octave_info": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int)]),
    "b200sift_detect_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "b200sift_detect_host_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "b200sift_detect_device_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200sift_detect_host_u8_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "b200sift_sync": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "b200sift_stream": (C.c_void_p, [C.c_void_p]),
    "b200sift_results": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                   C.POINTER(C.c_void_p), C.POINTER(C.c_int)]),
    "b200sift_copy_to_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "b200sift_debug_get_stack": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "b200sift_debug_get_kpt_buffer": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]),
    "b200sift_debug_get_counts": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int)]),
    "b200sift_debug_get_taps": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int]),
    "b200sift_debug_run_stages": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "b200sift_dump_frame": (C.c_int, [C.c_void_p, C.c_char_p]),
    "b200sift_rerun_stages": (C.c_int, [C.c_void_p, C.c_int]),
    "b200sift_frames_in_flight": (C.c_int, [C.c_void_p]),
    "b200sift_launches_per_frame": (C.c_int, [C.c_void_p]),
    "b200sift_time_frames": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float)]),
    "b200sift_time_frames_pipelined": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float)]),
    "b200sift_debug_timeline": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float)]),
    "b200sift_match_consecutive_async": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200sift_copy_results_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200sift_detect_batch_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200sift_time_stages": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float)]),
    "b200sift_profile_frame": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "b200sift_match_device": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "b200sift_match_host": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    "b200sift_match_device_noalloc": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "b200sift_match_noalloc_max_rows": (C.c_int, []),
    "b200sift_match_launches": (C.c_int, [C.c_int, C.c_int]),
    "b200sift_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "b200sift_comm_create": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "b200sift_comm_destroy": (C.c_int, [C.c_void_p]),
    "b200sift_comm_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b200sift_shard_rows": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "b200sift_match_sharded": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "b200sift_match_sharded_profile": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                                 C.POINTER(C.c_float)]),
    "b200sift_match_set_path": (C.c_int, [C.c_int]),
    "b200sift_debug_match_dist": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
}

_lib = None


def build_library(force: bool = False) -> str:
    """Compile the CUDA extension in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    args = ["make", "-C", CSRC, "-j8"]
    if force:
        subprocess.check_call(["make", "-C", CSRC, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(args, stdout=subprocess.DEVNULL)
    return LIB_PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200SiftError(
            f"{LIB_PATH} is missing: build it with `make -C {CSRC}` (or __graft_entry__.build()). "
            "The SIFT engine is CUDA-only; there is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the ABI and the header ever diverge
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc: int) -> int:
    if rc < 0:
        msg = load().b200sift_last_error()
        raise B200SiftError(f"b200sift error {rc}: {msg.decode() if msg else ''}")
    return rc
