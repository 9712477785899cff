"""OpenCV glue for the detector's results: the Python side of the reference's cvUtils
(cvUtils/Conversion.hh:12-71, Conversion.cc:22-58, ConversionImpl.hpp:7-82), same function names.
The C++ side is include/cvUtils/Conversion.hh. Host-side only; `cv2` is imported lazily.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np


def cvMatToImage(mat: np.ndarray, dtype=np.float32) -> np.ndarray:
    """ConversionImpl.hpp:7-32: single-channel uint8 / float32 / float64 image -> [rows, cols] array of
    `dtype`; values are cast, not scaled (8-bit input stays 0..255, which the thresholds assume)."""
    mat = np.asarray(mat)
    if mat.ndim == 3 and mat.shape[2] > 1 or mat.ndim > 3:
        raise RuntimeError("More channels than expected")  # ConversionImpl.hpp:9-11
    if mat.ndim == 3:
        mat = mat[:, :, 0]
    if mat.dtype not in (np.uint8, np.float32, np.float64):
        return np.zeros((0, 0), dtype)  # the reference leaves the image empty for other depths
    return np.ascontiguousarray(mat, dtype=dtype)


def imageToCvMat(image: np.ndarray) -> np.ndarray:
    """ConversionImpl.hpp:34-47: uint8 / float32 image -> CV_8UC1 / CV_32FC1 Mat (a deep copy)."""
    image = np.asarray(image)
    if image.dtype not in (np.uint8, np.float32):
        raise TypeError("8-bit or float images only")
    return np.array(image, copy=True, order="C")


def areEqual(a: np.ndarray, b: np.ndarray) -> bool:
    """Conversion.cc:10-20: same shape and bits."""
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and a.dtype == b.dtype and bool(np.array_equal(a.view(np.uint8), b.view(np.uint8)))


def normalize(image: np.ndarray) -> np.ndarray:
    """ConversionImpl.hpp:50-64: min-max stretch to 0..255 in the image's own type, truncating cast to uint8."""
    image = np.asarray(image)
    lo, hi = image.min(), image.max()
    scale = image.dtype.type(255) / (hi - lo)
    return ((image - lo) * scale).astype(np.uint8)


def localKptToCvKpt(kpts: np.ndarray, features: np.ndarray, size: int = -1) -> list:
    """Conversion.cc:22-45: pt = (x, y), octave = packed octave word, response = |contrast|, size, angle (deg).
    `kpts` [n,3] / `features` [n,4] are Detector.final_kpts / final_features; pass size=Detector.total_size."""
    import cv2

    assert len(kpts) == len(features)
    n = len(kpts) if size == -1 else size
    return [cv2.KeyPoint(x=float(kpts[i][0]), y=float(kpts[i][1]), size=float(features[i][1]), angle=float(features[i][3]),
                         response=float(features[i][2]), octave=int(features[i][0])) for i in range(n)]


def descriptorToCvMat(descriptors: np.ndarray, num_pts: int) -> np.ndarray:
    """ConversionImpl.hpp:66-82: [num_pts, 128] float32, row i = descriptor of keypoint i (half 0..512 widened)."""
    d = np.asarray(descriptors).reshape(-1, 128)
    return np.ascontiguousarray(d[: min(num_pts, len(d))], dtype=np.float32)


def cvtMatchToDMatch(match: Sequence[int]) -> list:
    """Conversion.cc:47-58: one DMatch(queryIdx=i, trainIdx=match[i], distance=0) per accepted match."""
    import cv2

    return [cv2.DMatch(int(i), int(m), 0.0) for i, m in enumerate(match) if m != -1]
