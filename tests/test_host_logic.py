"""Host-side logic that needs no GPU: octave geometry, synthetic input determinism, sharding, and the
world_size-2 gloo run of the sharded matcher plumbing."""
import hashlib
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT


@pytest.mark.parametrize("cols,rows,n_oct,dims", [
    (752, 480, 9, [(752, 480), (376, 240), (188, 120), (94, 60), (47, 30), (24, 15), (12, 8), (6, 4), (3, 2)]),
    (1600, 900, 10, [(1600, 900), (800, 450), (400, 225), (200, 113), (100, 56), (50, 28), (25, 14), (13, 7), (6, 4), (3, 2)]),
    (1920, 1200, 10, [(1920, 1200), (960, 600), (480, 300), (240, 150), (120, 75), (60, 38), (30, 19), (15, 9), (8, 5), (4, 2)]),
])
def test_octave_geometry_of_oracle(cols, rows, n_oct, dims):
    # Detector.hh:27 and Detector.cu:83
    from oracle.pyoracle import CpuOracle

    o = CpuOracle(cols, rows, 5000)
    assert o.n_octaves == n_oct
    assert [(x["w"], x["h"]) for x in o.octaves] == dims
    assert o.octaves[0]["cap"] == 50000


def test_synth_inputs_are_deterministic():
    from another_cuda_sift_b200.synth import next_frame, synth_descriptors, synth_frame

    a, b = synth_frame(120, 160, 0), synth_frame(120, 160, 0)
    assert a.dtype == np.float32 and np.array_equal(a, b)
    assert a.min() >= 0 and a.max() <= 255 and np.array_equal(a, np.floor(a))
    assert not np.array_equal(a, synth_frame(120, 160, 1))
    n = next_frame(a, 1)
    assert n.shape == a.shape and np.array_equal(n, next_frame(a, 1))
    q, t = synth_descriptors(64, 80)
    assert q.shape == (64, 128) and t.shape == (80, 128) and q.dtype == np.float16
    assert hashlib.sha1(q.tobytes()).hexdigest() == hashlib.sha1(synth_descriptors(64, 80)[0].tobytes()).hexdigest()
    assert float(q.astype(np.float32).max()) <= 512.0 and float(q.astype(np.float32).min()) >= 0.0


def test_frame_and_row_shards_partition_the_work():
    from another_cuda_sift_b200.sharding import frame_shard, row_shard

    for n in (0, 1, 7, 8, 256, 100000):
        for world in (1, 2, 4, 8):
            frames = [list(frame_shard(n, world, r)) for r in range(world)]
            assert sorted(sum(frames, [])) == list(range(n))
            rows = [row_shard(n, world, r) for r in range(world)]
            covered = sum((list(range(a, b)) for a, b in rows), []) if n < 1000 else None
            if covered is not None:
                assert covered == list(range(n))
            assert sum(b - a for a, b in rows) == n
            assert max(b - a for a, b in rows) <= -(-n // world) if n else True
    # overlap=1: every consecutive pair (i-1, i) is local to the rank that owns frame i
    for world in (2, 4, 8):
        for r in range(world):
            own = frame_shard(32, world, r)
            ext = frame_shard(32, world, r, overlap=1)
            assert ext.stop == own.stop and ext.start == max(0, own.start - 1)


def test_sharded_match_gloo_world2(tmp_path):
    """Two CPU processes over gloo: each matches its query-row block (compute = the CPU oracle, here only
    as a stand-in for the CUDA kernel), all-gather reassembles the same answer as one unsharded call."""
    script = os.path.join(ROOT, "tests", "_gloo_worker.py")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29617")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29617", script, str(tmp_path)],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()
