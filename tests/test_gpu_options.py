"""b200sift_options: frames in flight and the three opt-in fixes of the reference's own TODO list.
With every option off the engine is the reference (tests/test_gpu_parity.py pins that against goldens); with an option
on there is no reference implementation to compare with, so the oracle is the CPU restatement carrying the same fix
(oracle/sift_oracle.cpp, stated there) plus properties against the engine's own default output."""
import numpy as np
import pytest

from util import check_results

pytestmark = pytest.mark.gpu


def _det(cols, rows, nf=5000, **opts):
    from another_cuda_sift_b200 import CudaSiftConfig, Detector

    cfg = {k: opts.pop(k) for k in list(opts) if k in ("upscale", "numOctaveLayers", "sigma")}
    d = Detector(CudaSiftConfig(col_width=cols, row_width=rows, numFeatures=nf, **cfg), **opts)
    assert d.gpuWarmUpAndAllocate()
    return d


def _res(det):
    det.copyToHost(True)
    n = det.total_size
    return dict(n=n, kpts=det.final_kpts[:n].copy(), feats=det.final_features[:n].copy(), desc=det.descriptors[:n].copy())


def _frames(rows, cols, seed, n=2):
    from another_cuda_sift_b200.synth import next_frame, synth_frame

    f = [synth_frame(rows, cols, seed)]
    for i in range(1, n):
        f.append(next_frame(f[-1], seed + i))
    return f


def test_remove_duplicates():
    from oracle.pyoracle import CpuOracle

    cols, rows = 752, 480
    base, dd, cpu = _det(cols, rows), _det(cols, rows, remove_duplicates=True), CpuOracle(cols, rows, 5000, remove_duplicates=True)
    found = 0
    for img in _frames(rows, cols, 0, 3):
        base.detectAndCompute(img)
        dd.detectAndCompute(img)
        cpu.detect(img)
        a, b = _res(base), _res(dd)
        key = lambda r: np.concatenate([r["kpts"][:, :2], r["feats"][:, [1, 3]]], axis=1)
        ka, kb = key(a), key(b)
        # which octave LIST an output row came from: the output is the concatenation of the octaves' lists (T_o rows each).
        # Not the packed octave field: a never-written slot (quirk Q5) can be all zeros in several octaves' lists.
        T = [base.debug_counts(o)["T"] for o in range(base.n_octaves)]
        assert sum(T) == a["n"]
        lst = np.repeat(np.arange(base.n_octaves), T)
        # property: the survivors are exactly the default output minus later duplicates of (x, y, size, angle) within an
        # octave's list, order kept, descriptors untouched
        seen, keep = set(), []
        for i, k in enumerate(ka):
            t = (k[0], k[1], k[2], k[3], int(lst[i]))
            if t not in seen:
                seen.add(t)
                keep.append(i)
        found += len(ka) - len(keep)
        assert b["n"] == len(keep)
        assert np.array_equal(kb, ka[keep]) and np.array_equal(b["desc"].view(np.uint16), a["desc"][keep].view(np.uint16))
        Tb = [dd.debug_counts(o)["T"] for o in range(dd.n_octaves)]
        lb = np.repeat(np.arange(dd.n_octaves), Tb)
        assert len({tuple(k) + (int(o),) for k, o in zip(kb, lb)}) == len(kb)
        check_results(b, cpu.results(), min_agree=0.97, desc_tol=3e-2, desc_frac=0.95, size_rtol=1e-5)
    assert found > 0, "the test frames no longer contain duplicate keypoints: the check above is vacuous"


def test_interpolate_orientation():
    from oracle.pyoracle import CpuOracle

    cols, rows = 640, 360
    base, it, cpu = _det(cols, rows), _det(cols, rows, interpolate_orientation=True), CpuOracle(cols, rows, 5000, interpolate_orientation=True)
    for img in _frames(rows, cols, 5, 2):
        base.detectAndCompute(img)
        it.detectAndCompute(img)
        cpu.detect(img)
        a, b, c = _res(base), _res(it), cpu.results()
        # property: same keypoints in the same order; only the angle moves, by less than one histogram bin
        assert a["n"] == b["n"] and np.array_equal(a["kpts"], b["kpts"]) and np.array_equal(a["feats"][:, :3], b["feats"][:, :3])
        d = np.abs(a["feats"][:, 3] - b["feats"][:, 3])
        d = np.minimum(d, 360.0 - d)
        assert d.max() <= 10.0 and (d > 1e-3).mean() > 0.9
        assert (b["feats"][:, 3] >= 0).all() and (b["feats"][:, 3] < 360.0).all()   # (0, 360] becomes [0, 360)
        # vs the CPU restatement of the same fix: positions identical, angles within 1e-2 rad (the histograms differ by
        # MUFU-vs-libm rounding, the parabola amplifies that a little)
        n = min(b["n"], c["n"])
        same = np.all(b["kpts"][:n] == c["kpts"][:n], axis=1) & (b["feats"][:n, 0] == c["feats"][:n, 0])
        assert same.mean() >= 0.97
        da = np.abs(b["feats"][:n, 3] - c["feats"][:n, 3])[same]
        da = np.minimum(da, 360.0 - da)
        assert (da <= np.degrees(1e-2)).mean() >= 0.99, float(da.max())


def test_working_upscale():
    from another_cuda_sift_b200 import B200SiftError, CudaSiftConfig, Detector
    from oracle.pyoracle import CpuOracle

    cols, rows = 320, 200
    with pytest.raises(B200SiftError):   # still rejected without the opt-in (CudaSiftConfig.hh:11-12)
        Detector(CudaSiftConfig(col_width=cols, row_width=rows, upscale=True)).gpuWarmUpAndAllocate()
    up, cpu = _det(cols, rows, upscale=True, working_upscale=True), CpuOracle(cols, rows, 5000, working_upscale=True)
    assert (up.octaves[0]["w"], up.octaves[0]["h"]) == (2 * cols, 2 * rows)
    assert [(o["w"], o["h"], o["pitch"], o["cap"]) for o in up.octaves] == [(o["w"], o["h"], o["pitch"], o["cap"]) for o in cpu.octaves]
    base = _det(cols, rows)
    for img in _frames(rows, cols, 9, 2):
        cpu.detect(img)
        up.debug_run_stages(img, 4)
        for oi in range(up.n_octaves):
            assert np.array_equal(up.debug_gauss(oi).view(np.uint32), cpu.stack(oi, "gauss").view(np.uint32)), oi
            c, d = cpu.counts(oi), up.debug_counts(oi)
            assert (d["P"], d["M"]) == (c["P"], c["M"]), (oi, c, d)
        up.detectAndCompute(img)
        base.detectAndCompute(img)
        r = _res(up)
        check_results(r, cpu.results(), min_agree=0.97, desc_tol=3e-2, desc_frac=0.95, size_rtol=1e-5)
        # properties: positions are input pixels (inside the frame, half-pixel grid), octave -1 exists, and the doubled
        # frame finds more keypoints than the plain one
        assert r["n"] > _res(base)["n"]
        # (the last entry of every octave's list is the reference's never-written slot, quirk Q5: it can hold raw
        # coordinates of the doubled frame, so up to one entry per octave is exempt)
        outside = (r["kpts"][:, 0] >= cols) | (r["kpts"][:, 1] >= rows)
        assert outside.sum() <= up.n_octaves, int(outside.sum())
        assert np.array_equal(r["kpts"][:, :2] * 2, np.round(r["kpts"][:, :2] * 2))
        octv = r["feats"][:, 0].astype(np.int64) & 255
        assert (octv == 255).any() and (octv[octv != 255] < up.n_octaves).all()


@pytest.mark.parametrize("slots", [3, 6, 32])
def test_frames_in_flight_option(slots):
    import torch

    from another_cuda_sift_b200 import B200SiftError, CudaSiftConfig, Detector, HostResults

    cols, rows, nf = 640, 360, 2000
    frames = _frames(rows, cols, 21, 2 * slots + 1)
    ref = _det(cols, rows, nf)
    want = []
    for f in frames:
        ref.detectAndCompute(f)
        want.append(_res(ref))
    det = _det(cols, rows, nf, frames_in_flight=slots)
    assert det.frames_in_flight == slots
    res = [HostResults(nf) for _ in frames]
    midx = [torch.empty(nf, dtype=torch.int32).pin_memory() for _ in frames]
    det.detectAndComputeBatch([torch.from_numpy(f).pin_memory().numpy() for f in frames], res, midx)
    for w, r in zip(want, res):
        n = w["n"]
        assert r.total == n and np.array_equal(r.kpts[:n].numpy(), w["kpts"]) and np.array_equal(r.desc[:n].numpy().view(np.uint16), w["desc"].view(np.uint16))
    for bad in (1, 2, 33):
        with pytest.raises(B200SiftError):
            Detector(CudaSiftConfig(col_width=64, row_width=48), frames_in_flight=bad).gpuWarmUpAndAllocate()


def test_debug_timeline_is_ordered():
    """b200sift_debug_timeline (development aid behind tools/dev_timeline.py): for every frame the four time stamps
    (stream reaches part A, end of A, end of B, end of C) are ordered, part B ends in frame order (the shared candidate
    lists), and the replay leaves the detector usable with unchanged results."""
    import ctypes as C

    from another_cuda_sift_b200 import _lib
    from another_cuda_sift_b200.synth import synth_frame

    cols, rows = 320, 240
    det = _det(cols, rows, 2000, frames_in_flight=4)
    img = synth_frame(rows, cols, 9)
    for _ in range(3):   # the same frame repeatedly: the stale-slot history (quirk Q5) has settled after the second
        det.detectAndCompute(img)
    det.copyToHost(True)
    n0, d0 = det.total_size, det.descriptors[:det.total_size].copy()
    assert n0 > 50
    n = 23
    out = (C.c_float * (4 * n))()
    _lib.check(det.lib.b200sift_debug_timeline(det._h, n, out))
    t = np.array(list(out)).reshape(n, 4)
    assert (np.diff(t, axis=1) >= 0).all(), t
    assert (np.diff(t[:, 2]) > 0).all(), t[:, 2]
    assert t.max() < 1e3
    det.detectAndCompute(img)
    det.copyToHost(True)
    assert det.total_size == n0 and np.array_equal(det.descriptors[:n0].view(np.uint16), d0.view(np.uint16))
