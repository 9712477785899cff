"""B200-native (sm_100a) drop-in for the data-parallel hot path of Cryptum169/another-cuda-sift.

Public surface mirrors the reference: CudaSiftConfig, Detector (gpuWarmUpAndAllocate /
detectAndCompute / copyToHost) and matchBruteForce; the C ABI lives in include/b200sift.h.
"""
from ._lib import B200SiftError, build_library, load  # noqa: F401
from .detector import CudaSiftConfig, Detector, DeviceArray, HostResults, matchBruteForce  # noqa: F401

__all__ = ["CudaSiftConfig", "Detector", "DeviceArray", "HostResults", "matchBruteForce", "B200SiftError", "build_library", "load"]
