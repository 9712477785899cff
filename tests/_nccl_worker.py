"""world_size-N NCCL worker for tests/test_gpu_multi.py::test_sharded_matcher_nccl: the native
b200sift_match_sharded (ncclBroadcast of the train set + matcher on this rank's rows + ncclAllGather) against the
single-GPU matcher and the CPU oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from another_cuda_sift_b200 import matchBruteForce  # noqa: E402
from another_cuda_sift_b200.sharding import ShardedMatcher  # noqa: E402
from another_cuda_sift_b200.synth import synth_descriptors  # noqa: E402
from oracle.pyoracle import CpuOracle  # noqa: E402  (test infrastructure)


def main():
    out = sys.argv[1]
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    sm = ShardedMatcher()
    assert (sm.rank, sm.world) == (rank, world)
    cpu = CpuOracle(64, 64, 100)
    # (nq, nt): resident-chunk filter, streaming filter (> 8192 train rows), ragged blocks (nq not divisible), tiny
    for nq, nt in ((1000, 3000), (2051, 9000), (5, 300), (3, 20000)):
        q, t = synth_descriptors(max(nq, 8), max(nt, 8))
        q, t = q[:nq], t[:nt]
        lo, hi = sm.rows(nq)
        dq = torch.from_numpy(q[lo:hi].copy().view(np.int16)).cuda() if hi > lo else None
        # the train set exists on rank 0 only; the others pass a zeroed receive buffer
        dt = torch.from_numpy(t.view(np.int16)).cuda() if rank == 0 else torch.zeros((nt, 128), dtype=torch.int16, device="cuda")
        got = sm.match(dq, nq, dt, nt, train_root=0)
        torch.cuda.synchronize()
        got = got.cpu().numpy()
        assert np.array_equal(dt.cpu().numpy().view(np.float16), t), "broadcast did not deliver the train set"
        full_q = torch.from_numpy(q.view(np.int16)).cuda()
        want = matchBruteForce(full_q, nq, dt, nt, stream=torch.cuda.current_stream().cuda_stream).cpu().numpy()
        assert np.array_equal(got, want), (rank, nq, nt)
        rows = np.random.default_rng(rank).choice(nq, size=min(16, nq), replace=False)
        assert np.array_equal(cpu.match(q[rows], t), got[rows]), (rank, nq, nt)
        # train set already resident everywhere (train_root = -1): same answer, no broadcast
        again = sm.match(dq, nq, dt, nt, train_root=-1)
        torch.cuda.synchronize()
        assert np.array_equal(again.cpu().numpy(), want)
    prof = sm.profile(dq, nq, dt, nt, train_root=0)
    assert set(prof) == {"broadcast", "norms", "filter", "rerank", "allgather"} and all(v >= 0 for v in prof.values())
    sm.close()
    open(os.path.join(out, f"ok{rank}"), "w").write("ok")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
