"""Pins the CPU oracle (oracle/sift_oracle.cpp) against golden vectors captured from the UNMODIFIED
reference CUDA build on a B200 (tests/golden/make_golden.py). CPU-only.

Bars: taps, Gaussian pyramid, raw extrema, refined keypoints, matcher indices: bit-exact.
Orientation / descriptor: tolerance (the reference uses MUFU approximations and float atomics)."""
import numpy as np
import pytest

from conftest import golden, golden_frames
from util import check_results


def test_taps_bit_exact():
    from oracle.pyoracle import CpuOracle

    g = golden("taps.npz")
    o = CpuOracle(64, 64, 100)
    for i in range(-1, 6):
        assert np.array_equal(o.taps(i).view(np.uint32), g[f"taps{i}"].view(np.uint32)), i
    assert [len(g[f"taps{i}"]) for i in range(-1, 6)] == [9, 11, 9, 11, 13, 17, 21]


@pytest.mark.parametrize("name", ["frames_160x120.npz", "frames_203x151.npz"])
def test_small_frames_stage_by_stage(name):
    from oracle.pyoracle import CpuOracle

    g = golden(name)
    frames = golden_frames(g)
    o = CpuOracle(int(g["cols"]), int(g["rows"]), int(g["num_features"]))
    assert o.n_octaves == int(g["n_octaves"])
    assert np.array_equal(np.array([[x["w"], x["h"], x["pitch"], x["cap"]] for x in o.octaves]), g["octaves"])
    for fr in range(int(g["nframes"])):
        o.set_image(frames[fr])
        for s in range(8):
            o.stage(s)
        counts = g[f"f{fr}_counts"]
        for oi in range(o.n_octaves):
            if f"f{fr}_gauss{oi}" in g:
                assert np.array_equal(o.stack(oi, "gauss").view(np.uint32), g[f"f{fr}_gauss{oi}"].view(np.uint32)), (fr, oi)
            c = o.counts(oi)
            assert c["P"] == counts[0][oi] and c["M"] == counts[1][oi], (fr, oi, c, counts[:, oi])
        want = dict(n=int(g[f"f{fr}_total"]), kpts=g[f"f{fr}_kpts"], feats=g[f"f{fr}_feats"], desc=g[f"f{fr}_desc"])
        check_results(o.results(), want, min_agree=0.97, desc_tol=3e-2, desc_frac=0.95, size_rtol=1e-5, octaves=g["octaves"])


@pytest.mark.parametrize("name", ["frames_752x480.npz", "frames_caps_752x480.npz"])
def test_752x480_history(name):
    from oracle.pyoracle import CpuOracle

    g = golden(name)
    frames = golden_frames(g)
    o = CpuOracle(752, 480, int(g["num_features"]))
    for fr in range(int(g["nframes"])):
        o.set_image(frames[fr])
        for s in range(4):
            o.stage(s)
        for oi in range(o.n_octaves):
            sums = o.stack(oi, "gauss").astype(np.float64).sum(axis=(1, 2))
            xors = np.bitwise_xor.reduce(o.stack(oi, "gauss").view(np.uint32).reshape(6, -1), axis=1)
            assert np.array_equal(sums, g[f"f{fr}_gauss_sum"][oi]) and np.array_equal(xors, g[f"f{fr}_gauss_xor"][oi])
            ext = g[f"f{fr}_extrema{oi}"]
            assert np.array_equal(o.kpt_buffer(oi, "cand")[:len(ext)].astype(np.int16), ext), (fr, oi)
        for s in range(4, 8):
            o.stage(s)
        counts = g[f"f{fr}_counts"]
        for oi in range(o.n_octaves):
            c = o.counts(oi)
            assert (c["P"], c["M"]) == (counts[0][oi], counts[1][oi])
            assert abs(c["T"] - counts[2][oi]) <= max(2, 0.03 * counts[2][oi])
        want = dict(n=int(g[f"f{fr}_total"]), kpts=g[f"f{fr}_kpts"], feats=g[f"f{fr}_feats"], desc=g[f"f{fr}_desc"])
        check_results(o.results(), want, min_agree=0.97, desc_tol=3e-2, desc_frac=0.95, size_rtol=1e-5, octaves=g["octaves"])


def test_refinement_bit_exact_752():
    from oracle.pyoracle import CpuOracle

    g = golden("frames_752x480.npz")
    frames = golden_frames(g)
    o = CpuOracle(752, 480, 5000)
    o.set_image(frames[0])
    for s in range(5):
        o.stage(s)
    for oi in range(o.n_octaves):
        want, wantF = g[f"f0_refined{oi}"], g[f"f0_refinedF{oi}"]
        n = len(want)
        assert np.array_equal(o.kpt_buffer(oi, "coll")[:n], want)
        gotF = o.kpt_buffer(oi, "collF")[:n, :3]
        assert np.array_equal(gotF[:, 0], wantF[:, 0])
        assert np.array_equal(gotF[:, 2], wantF[:, 2])  # |contrast|: same FMA chain
        if n:  # size goes through powf: host libm vs CUDA libdevice, <= 2 ulp
            assert np.max(np.abs(gotF[:, 1] - wantF[:, 1]) / wantF[:, 1]) < 3e-7


def test_matcher_bit_exact():
    from another_cuda_sift_b200.synth import synth_descriptors
    from oracle.pyoracle import CpuOracle

    g = golden("match.npz")
    q, t = synth_descriptors(2000, 2000)
    o = CpuOracle(64, 64, 100)
    assert np.array_equal(o.match(q[:400], t), g["m2000"][:400])
    for nq, nt in ((5, 1), (7, 2), (64, 33), (100, 31), (33, 64), (257, 1000)):
        assert np.array_equal(o.match(q[:nq], t[:nt]), g[f"m{nq}x{nt}"]), (nq, nt)
    assert np.array_equal(o.match(q[:300], q[:300]), g["self300"])
