"""cvUtils converters (reference cvUtils/Conversion.hh:12-71): the C++ header against a
minimal OpenCV stand-in, the Python side against the real cv2. Host-side only."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT


def test_cpp_header_compiles_and_converts(tmp_path):
    exe = str(tmp_path / "cvutils_check")
    subprocess.check_call(["g++", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tests", "shim"),
                           "-I", "/usr/local/cuda/include", os.path.join(ROOT, "tests", "cvutils_check.cc"), "-o", exe])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0 and "ok" in out.stdout, out.stdout + out.stderr


def test_header_keeps_reference_names():
    src = open(os.path.join(ROOT, "include", "cvUtils", "Conversion.hh")).read()
    for name in ("namespace OpencvUtils", "cvMatToImage", "imageToCvMat", "areEqual", "normalize", "localKptToCvKpt",
                 "descriptorToCvMat", "cvtMatchToDMatch"):
        assert name in src, name


def test_python_converters_with_cv2():
    cv2 = pytest.importorskip("cv2")
    from another_cuda_sift_b200 import cvutils

    img8 = (np.arange(12, dtype=np.uint8) * 20).reshape(3, 4)
    f = cvutils.cvMatToImage(img8)
    assert f.dtype == np.float32 and f[2, 3] == 220.0            # cast, not scaled
    with pytest.raises(RuntimeError):
        cvutils.cvMatToImage(np.zeros((3, 4, 3), np.uint8))
    assert cvutils.areEqual(cvutils.imageToCvMat(f), f) and not cvutils.areEqual(f, img8)
    n8 = cvutils.normalize(f)
    assert n8.dtype == np.uint8 and n8.min() == 0 and n8.max() == 255

    kpts = np.zeros((3, 3), np.float32)
    feats = np.zeros((3, 4), np.float32)
    kpts[1] = (10, 20, 2)
    feats[1] = (1 + (2 << 8) + (255 << 16), 3.5, 0.05, 270)
    k = cvutils.localKptToCvKpt(kpts, feats, 2)
    assert len(k) == 2 and isinstance(k[1], cv2.KeyPoint)
    assert k[1].pt == (10.0, 20.0) and k[1].octave == 1 + (2 << 8) + (255 << 16)
    assert (k[1].size, k[1].angle) == (3.5, 270.0) and abs(k[1].response - 0.05) < 1e-7
    assert len(cvutils.localKptToCvKpt(kpts, feats)) == 3

    desc = np.arange(256, dtype=np.float16).reshape(2, 128)
    m = cvutils.descriptorToCvMat(desc, 5)
    assert m.shape == (2, 128) and m.dtype == np.float32 and m[1, 127] == 255.0

    dm = cvutils.cvtMatchToDMatch([4, -1, 0, -1, 2])
    assert [(d.queryIdx, d.trainIdx) for d in dm] == [(0, 4), (2, 0), (4, 2)] and all(d.distance == 0 for d in dm)
    # the converted types are what OpenCV pipelines consume: a BFMatcher over the float descriptors accepts them
    bf = cv2.BFMatcher(cv2.NORM_L2)
    assert len(bf.match(m, m)) == 2
