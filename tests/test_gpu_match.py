"""Matcher parity through the C ABI: bit-exact indices vs golden (reference on B200), vs the CPU oracle
(which restates Match.cu:35-177 exactly, incl. the Q15 merge), edge cases, and size-independent
properties at sizes the CPU oracle cannot finish."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


def test_2000x2000_vs_golden(path):
    from another_cuda_sift_b200 import matchBruteForce
    from another_cuda_sift_b200.synth import synth_descriptors

    g = golden("match.npz")
    q, t = synth_descriptors(2000, 2000)
    got = matchBruteForce(q, 2000, t, 2000)
    assert np.array_equal(got, g["m2000"])
    assert (got >= 0).sum() > 200
    for nq, nt in ((5, 1), (7, 2), (64, 33), (100, 31), (33, 64), (257, 1000)):
        assert np.array_equal(matchBruteForce(q[:nq], nq, t[:nt], nt), g[f"m{nq}x{nt}"]), (nq, nt)
    assert np.array_equal(matchBruteForce(q[:300], 300, q[:300], 300), g["self300"])


@pytest.fixture(params=[0, 1, 2, 3], ids=["tensor-core", "cuda-core", "tensor-core-unpipelined", "tensor-core-single-pass-stream"])
def path(request):
    """Both matcher implementations (csrc/match_tc.cu: tcgen05 filter + exact re-rank; csrc/match.cu: exact
    CUDA-core kernel) must reproduce the reference bit for bit."""
    from another_cuda_sift_b200 import _lib

    lib = _lib.load()
    lib.b200sift_match_set_path(request.param)
    yield request.param
    lib.b200sift_match_set_path(0)


# (200, 5000): one 256-wide TMEM-resident tile per chunk; (150, 9000) ... : streamed 128-row tiles, two passes
# (8193 -> per 257: three tiles, the last one with a single row; 12289 -> per 385: the last chunk is short)
@pytest.mark.parametrize("nq,nt", [(1, 1), (3, 2), (31, 32), (33, 31), (100, 257), (513, 700), (1, 4000), (700, 5),
                                   (200, 5000), (150, 9000), (300, 8193), (129, 12289), (257, 16500)])
def test_vs_cpu_oracle(nq, nt, path):
    from another_cuda_sift_b200 import matchBruteForce
    from another_cuda_sift_b200.synth import synth_descriptors
    from oracle.pyoracle import CpuOracle

    q, t = synth_descriptors(max(nq, 8), max(nt, 8))
    q, t = q[:nq], t[:nt]
    assert np.array_equal(matchBruteForce(q, nq, t, nt), CpuOracle(64, 64, 100).match(q, t))


def test_paths_agree_on_detector_descriptors(path):
    """Real descriptors (consecutive frames, as tool/extract_and_match_example.cc:82 uses the matcher)."""
    from another_cuda_sift_b200 import CudaSiftConfig, Detector, matchBruteForce
    from another_cuda_sift_b200.synth import next_frame, synth_frame
    from oracle.pyoracle import CpuOracle

    det = Detector(CudaSiftConfig(col_width=752, row_width=480))
    det.gpuWarmUpAndAllocate()
    a = synth_frame(480, 752, 21)
    det.detectAndCompute(a)
    det.copyToHost(True)
    n0, d0 = det.total_size, det.descriptors[:det.total_size].copy()
    det.detectAndCompute(next_frame(a, 22))
    det.copyToHost(True)
    n1, d1 = det.total_size, det.descriptors[:det.total_size].copy()
    got = matchBruteForce(d0, n0, d1, n1)
    assert np.array_equal(got, CpuOracle(64, 64, 100).match(d0, d1))
    assert (got >= 0).mean() > 0.2


def test_adversarial_ties_and_duplicates_vs_cpu_oracle(path):
    """Duplicate train rows (exact ties: lowest index must win, strict '<'), near ties at the ratio
    threshold, all-zero rows, saturated values."""
    from another_cuda_sift_b200 import matchBruteForce
    from oracle.pyoracle import CpuOracle

    rng = np.random.default_rng(9)
    base = (rng.random((64, 128)) * 120).astype(np.float16)
    t = np.concatenate([base, base, base[::-1], np.zeros((3, 128), np.float16), np.full((2, 128), 512, np.float16)])
    q = np.concatenate([base[:40], (base[:40].astype(np.float32) + rng.integers(-2, 3, (40, 128))).clip(0, 512).astype(np.float16),
                        np.zeros((2, 128), np.float16)])
    got = matchBruteForce(q, len(q), t, len(t))
    assert np.array_equal(got, CpuOracle(64, 64, 100).match(q, t))


def test_streaming_filter_adversarial_vs_cpu_oracle(path):
    """Large train sets (> 8192 rows: the streamed two-pass tcgen05 filter). (a) a small base repeated many times:
    exact ties inside and across tiles and chunks, survivor lists overflow and fall back to whole-chunk scoring;
    (b) rows whose squared norm exceeds what the three-half split of |b|^2 can carry (> 2e9): the filter hands
    every chunk to the exact path; the half arithmetic of the reference overflows to inf there and so must ours."""
    from another_cuda_sift_b200 import matchBruteForce
    from oracle.pyoracle import CpuOracle

    rng = np.random.default_rng(17)
    base = (rng.random((300, 128)) * 200).astype(np.float16)
    t = np.concatenate([base] * 30 + [base[:77]])                      # 9077 rows, per = 284
    t[5000:5300] = (t[5000:5300].astype(np.float32) + rng.integers(-1, 2, (300, 128))).clip(0, 512).astype(np.float16)
    q = np.concatenate([base[:48], (base[100:148].astype(np.float32) + rng.integers(-2, 3, (48, 128))).clip(0, 512).astype(np.float16)])
    cpu = CpuOracle(64, 64, 100)
    assert np.array_equal(matchBruteForce(q, len(q), t, len(t)), cpu.match(q, t))
    big = (rng.random((8500, 128)) * 6000).astype(np.float16)         # |b|^2 ~ 1.5e9 .. 4e9
    big[::7] = (rng.random((len(big[::7]), 128)) * 100).astype(np.float16)
    qb = np.concatenate([big[:20], (rng.random((20, 128)) * 100).astype(np.float16)])
    assert np.array_equal(matchBruteForce(qb, len(qb), big, len(big)), cpu.match(qb, big))


def test_overflowed_survivor_lists_rescore_whole_chunks(path):
    """More than 7 survivors in one (query, chunk) record: the re-rank kernel re-scores the whole chunk, 16 columns per
    step. A query's best match (an exact or almost exact copy) sits among 11 noisy copies that all pass the filter's
    slack; chunk lengths that are not multiples of 16, the best column at the very end of a chunk / of the train set /
    inside the last partial step, two overflowed chunks for one query; resident filter (nt <= 8192) and streaming
    filter (nt > 8192)."""
    from another_cuda_sift_b200 import matchBruteForce
    from oracle.pyoracle import CpuOracle

    cpu = CpuOracle(64, 64, 100)
    rng = np.random.default_rng(23)

    def noisy(row, n, dims=128):
        out = np.repeat(row[None].astype(np.float32), n, axis=0)
        out[:, :dims] += rng.integers(-1, 2, (n, dims))
        return out.clip(0, 512).astype(np.float16)

    for nt in (32 * 37 + 5, 8192 + 32 * 11 + 3):
        per = -(-nt // 32)
        t = (rng.random((nt, 128)) * 150).astype(np.float16)
        q = (20 + rng.random((24, 128)) * 110).astype(np.float16)
        for k in range(12):            # 11 noisy copies + the best one in the LAST column of chunk k (k even) / first of the 12 (k odd)
            hi = min(nt, (k + 1) * per)
            t[hi - 12:hi] = noisy(q[k], 12)
            t[hi - 1 if k % 2 == 0 else hi - 12] = q[k] if k % 4 < 2 else noisy(q[k], 1, dims=6)[0]
        for k in range(12, 18):        # spread over a whole chunk, best one anywhere
            rows = (k + 1) * per + rng.permutation(per)[:10]
            t[rows] = noisy(q[k], 10)
            t[rows[k % 10]] = noisy(q[k], 1, dims=4)[0]
        t[nt - 9:] = noisy(q[18], 9)   # the last (short) chunk ends with 9 copies, the exact one last
        t[nt - 1] = q[18]
        t[5 * per + 3:5 * per + 14] = noisy(q[19], 11)   # query 19: two overflowed chunks, the best in the second
        t[9 * per:9 * per + 11] = noisy(q[19], 11)
        t[9 * per + 10] = q[19]
        want = cpu.match(q, t)
        assert (want[:20] >= 0).sum() >= 15, want   # the construction really produces matches, not a column of -1
        got = matchBruteForce(q, len(q), t, nt)
        assert np.array_equal(got, want), nt


def test_empty_inputs():
    from another_cuda_sift_b200 import matchBruteForce

    q = np.zeros((4, 128), np.float16)
    assert matchBruteForce(q, 0, q, 4).shape == (0,)
    # no train rows: sentinels only (Match.cu:125-129) -> 1e6 < 0.8e6 is false -> -1 everywhere
    assert np.array_equal(matchBruteForce(q, 4, q, 0), np.full(4, -1, np.int32))


def test_device_api_and_large_properties():
    """20k x 20k on the device API: (a) a permuted exact copy of the query set matches itself wherever a
    match is accepted; (b) row blocks computed separately equal the full result (the sharding unit of
    SURVEY.md 8(e)); (c) a 2k x 2k slice equals the CPU oracle on sampled rows."""
    import torch

    from another_cuda_sift_b200 import matchBruteForce
    from another_cuda_sift_b200.sharding import row_shard
    from another_cuda_sift_b200.synth import synth_descriptors
    from oracle.pyoracle import CpuOracle

    n = 20000
    q, t = synth_descriptors(n, n)
    perm = np.random.default_rng(1).permutation(n)
    dq = torch.from_numpy(q.view(np.int16)).cuda()
    dperm = torch.from_numpy(q[perm].copy().view(np.int16)).cuda()
    dt = torch.from_numpy(t.view(np.int16)).cuda()
    s = torch.cuda.current_stream().cuda_stream
    self_idx = matchBruteForce(dq, n, dperm, n, stream=s).cpu().numpy()
    ok = self_idx >= 0
    assert ok.mean() > 0.9
    # a zero-distance train row exists for every query; whichever index wins has distance 0 == identical row
    assert np.array_equal(q[perm][self_idx[ok]].view(np.uint16), q[ok].view(np.uint16))
    full = matchBruteForce(dq, n, dt, n, stream=s).cpu().numpy()
    parts = []
    for r in range(8):
        lo, hi = row_shard(n, 8, r)
        parts.append(matchBruteForce(dq[lo:hi], hi - lo, dt, n, stream=s).cpu().numpy())
    assert np.array_equal(np.concatenate(parts), full)
    rows = np.random.default_rng(2).choice(n, 24, replace=False)
    assert np.array_equal(CpuOracle(64, 64, 100).match(q[rows], t), full[rows])
