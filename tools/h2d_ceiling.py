#!/usr/bin/env python
"""Host->device (and device->host) copy ceiling of the box, one process per GPU: a bare loop of pinned
cudaMemcpyAsync transfers of frame-sized buffers (1920x1200 fp32 = 9.2 MB), nothing else running. This is the
number the engine's end-to-end float path (`e2e` in bench.py: 9.2 MB in + 1.4 MB out per frame) is bounded by.

    python tools/h2d_ceiling.py                     # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/h2d_ceiling.py

Prints one JSON line (rank 0): per-GPU and aggregate GB/s for H2D alone, D2H alone and both directions at once,
max over ranks of the CUDA-event time between two barriers."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    affinity = "off"
    if "--affinity" in sys.argv:
        from bench import bind_to_gpu_numa

        affinity = bind_to_gpu_numa(local)
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbuf, nbytes, iters = 8, 1920 * 1200 * 4, 40
    host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(nbuf)]
    dev = [torch.empty(nbytes, dtype=torch.uint8, device="cuda") for _ in range(nbuf)]
    hback = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(nbuf)]
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(h2d, d2h):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        for timed in (False, True):
            barrier()
            e0.record()
            s_in.wait_event(e0)
            s_out.wait_event(e0)
            for i in range(iters if timed else 4):
                for b in range(nbuf):
                    if h2d:
                        with torch.cuda.stream(s_in):
                            dev[b].copy_(host[b], non_blocking=True)
                    if d2h:
                        with torch.cuda.stream(s_out):
                            hback[b].copy_(dev[b], non_blocking=True)
            e1.record(s_in)
            e2.record(s_out)
            barrier()
        ms = max(e0.elapsed_time(e1), e0.elapsed_time(e2))
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return iters * nbuf * nbytes / (ms * 1e-3) / 1e9  # GB/s per GPU and direction

    h = run(True, False)
    d = run(False, True)
    both = run(True, True)
    if rank == 0:
        print(json.dumps({"n_gpus": world, "buffer_bytes": nbytes, "cpu_affinity": affinity,
                          "h2d_GBs_per_gpu": round(h, 2), "h2d_GBs_aggregate": round(h * world, 2),
                          "d2h_GBs_per_gpu": round(d, 2), "d2h_GBs_aggregate": round(d * world, 2),
                          "bidir_GBs_per_gpu_per_direction": round(both, 2),
                          "frames_per_s_ceiling_f32_1920x1200": round(h * world * 1e9 / nbytes, 1)}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
