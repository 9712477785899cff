"""Multi-GPU / multi-thread paths of the C ABI: the NCCL query-row sharded matcher (csrc/shard.cu) and one process
driving several GPUs from several threads. The single-rank cases run on any box; the others need >= 2 visible devices
and are skipped otherwise (`gpurun --gpus 2`)."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_devices():
    import torch

    return torch.cuda.device_count()


def test_sharded_matcher_single_rank():
    """world = 1 through the same native call (a one-rank NCCL communicator): equals matchBruteForce and the oracle."""
    import torch

    from another_cuda_sift_b200 import matchBruteForce
    from another_cuda_sift_b200.sharding import ShardedMatcher
    from another_cuda_sift_b200.synth import synth_descriptors
    from oracle.pyoracle import CpuOracle

    sm = ShardedMatcher()
    assert (sm.rank, sm.world) == (0, 1)
    for nq, nt in ((700, 2500), (1500, 9100)):
        q, t = synth_descriptors(nq, nt)
        dq, dt = torch.from_numpy(q.view(np.int16)).cuda(), torch.from_numpy(t.view(np.int16)).cuda()
        got = sm.match(dq, nq, dt, nt).cpu().numpy()
        want = matchBruteForce(dq, nq, dt, nt, stream=torch.cuda.current_stream().cuda_stream).cpu().numpy()
        assert np.array_equal(got, want)
        rows = np.random.default_rng(3).choice(nq, 24, replace=False)
        assert np.array_equal(CpuOracle(64, 64, 100).match(q[rows], t), got[rows])
    prof = sm.profile(dq, nq, dt, nt)
    assert prof["filter"] > 0 and prof["rerank"] > 0
    sm.close()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_matcher_nccl(world, tmp_path):
    if _n_devices() < world:
        pytest.skip(f"needs {world} GPUs")
    script = os.path.join(ROOT, "tests", "_nccl_worker.py")
    port = str(29700 + world)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=port)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", port, script, str(tmp_path)],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert all((tmp_path / f"ok{i}").exists() for i in range(world))


def _thread_job(dev, frames, q, t, out, errors):
    try:
        import torch

        from another_cuda_sift_b200 import CudaSiftConfig, Detector, matchBruteForce

        torch.cuda.set_device(dev)   # cudaSetDevice is per thread: everything below runs on `dev`
        det = Detector(CudaSiftConfig(col_width=frames[0].shape[1], row_width=frames[0].shape[0], numFeatures=2000))
        assert det.gpuWarmUpAndAllocate()
        res = []
        for f in frames:
            det.detectAndCompute(f)
            det.copyToHost(True)
            n = det.total_size
            res.append((n, det.final_kpts[:n].copy(), det.final_features[:n].copy(), det.descriptors[:n].copy()))
        # consecutive-frame matcher (resident-chunk tcgen05 filter) and a large train set (streaming filter, TMA):
        # both set per-device kernel attributes on first use
        m_small = matchBruteForce(res[0][3], res[0][0], res[1][3], res[1][0])
        dq, dt = torch.from_numpy(q.view(np.int16)).cuda(), torch.from_numpy(t.view(np.int16)).cuda()
        m_large = matchBruteForce(dq, len(q), dt, len(t), stream=torch.cuda.current_stream().cuda_stream).cpu().numpy()
        out[dev] = (res, m_small, m_large)
        det.close()
    except Exception as e:  # noqa: BLE001
        import traceback

        errors.append((dev, traceback.format_exc()))


def test_threads_one_process(tmp_path):
    """One process, one host thread per GPU: every thread creates its own detector on its own
    device and uses both matcher paths concurrently; results are identical on every device. With a single visible GPU the
    two threads share device 0 (still exercises concurrent handles and the per-device scratch serialisation)."""
    from another_cuda_sift_b200.synth import next_frame, synth_descriptors, synth_frame

    ndev = min(_n_devices(), 4)
    devs = list(range(ndev)) if ndev >= 2 else [0, 0]
    frames = [synth_frame(360, 640, 71)]
    frames.append(next_frame(frames[0], 72))
    q, t = synth_descriptors(600, 9000)
    out, errors = {}, []
    if ndev >= 2:
        threads = [threading.Thread(target=_thread_job, args=(d, frames, q, t, out, errors)) for d in devs]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        assert not errors, errors
        base = out[0]
        for d in devs[1:]:
            for (n, k, f, de), (n0, k0, f0, d0) in zip(out[d][0], base[0]):
                assert n == n0 and np.array_equal(k, k0) and np.array_equal(f, f0) and np.array_equal(de.view(np.uint16), d0.view(np.uint16))
            assert np.array_equal(out[d][1], base[1]) and np.array_equal(out[d][2], base[2])
    else:
        outs = [{}, {}]
        threads = [threading.Thread(target=_thread_job, args=(0, frames, q, t, outs[i], errors)) for i in range(2)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        assert not errors, errors
        a, b = outs[0][0], outs[1][0]
        for (n, k, f, de), (n0, k0, f0, d0) in zip(a[0], b[0]):
            assert n == n0 and np.array_equal(k, k0) and np.array_equal(de.view(np.uint16), d0.view(np.uint16))
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
