"""B200-native (sm_100a) drop-in for the data-parallel hot path of Cryptum169/another-cuda-sift.

Public surface mirrors the reference: CudaSiftConfig, Detector (gpuWarmUpAndAllocate /
detectAndCompute / copyToHost) and matchBruteForce; the C ABI lives in include/b200sift.h.
"""
import os as _os

# 32 hardware work queues instead of CUDA's default 8: the engine keeps up to 32 frames in flight, each on its own stream
# (csrc/engine.cu: b200sift_on_load; profiles/r2_timeline.md). Read when the CUDA context is created, hence set at import,
# before the first CUDA call of the process; a value chosen by the user is kept.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from ._lib import B200SiftError, build_library, load  # noqa: F401
from .detector import CudaSiftConfig, Detector, DeviceArray, HostResults, matchBruteForce  # noqa: F401

__all__ = ["CudaSiftConfig", "Detector", "DeviceArray", "HostResults", "matchBruteForce", "B200SiftError", "build_library", "load"]
