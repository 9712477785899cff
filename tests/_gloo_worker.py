"""world_size-2 gloo worker for tests/test_host_logic.py::test_sharded_match_gloo_world2."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from another_cuda_sift_b200.sharding import frame_shard, sharded_match  # noqa: E402
from another_cuda_sift_b200.synth import synth_descriptors  # noqa: E402
from oracle.pyoracle import CpuOracle  # noqa: E402  (test infrastructure)


def main():
    out = sys.argv[1]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    assert world == 2
    q, t = synth_descriptors(101, 77)
    cpu = CpuOracle(64, 64, 100)

    def compute(d, n, s, m):
        return torch.from_numpy(cpu.match(d.numpy().view(np.float16), s.numpy().view(np.float16)))

    tq, tt = torch.from_numpy(q.view(np.int16)), torch.from_numpy(t.view(np.int16))
    got = sharded_match(tq, 101, tt, 77, compute=compute)
    want = cpu.match(q, t)
    assert np.array_equal(got.numpy(), want), (rank, got[:10], want[:10])
    # frames: disjoint cover, counts all-reduced like bench.py does
    mine = torch.tensor([len(frame_shard(33, world, rank))])
    dist.all_reduce(mine)
    assert int(mine) == 33
    open(os.path.join(out, f"ok{rank}"), "w").write("ok")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
