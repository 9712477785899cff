"""Matcher benchmark: 2k x 2k ratio-test latency and the query-row sharded large problem.

    python tools/bench_match.py                                   # 2k x 2k latency + 20k x 20k on one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/bench_match.py --large 100000                       # 100k x 100k, query rows sharded over N GPUs

Prints one JSON line (rank 0). Times are CUDA events on the launching stream, after warm-up, max over ranks;
the large problem's timed region contains the ncclBroadcast of the train set, the matcher on this rank's row block AND
the ncclAllGather of the int32 indices (native call b200sift_match_sharded). The descriptor sets are the seeded synthetic ones of SURVEY.md 8(d) (seed 42 / 43).
Roofline of the matcher: tensor-bound, flops = 2*nq*nt*128 against MEASURED_PEAKS.json's dense bf16 figure.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def tensor_peak_tflops():
    """Dense bf16 tensor throughput measured on this pool's B200s (driver-written MEASURED_PEAKS.json,
    sustained figure), else the profiling recipe's nominal 2250."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "measured"
    except Exception:  # noqa: BLE001
        return 2250.0, "nominal"


def match_latency(n: int, iters: int = 200, warmup: int = 20, rotate: int = 8):
    """Device-resident n x n latency (median of per-call CUDA-event times, `rotate` distinct descriptor
    sets so successive calls do not hit the same L2 lines) and the blocking host-buffer call."""
    import time

    import torch

    from another_cuda_sift_b200 import matchBruteForce
    from another_cuda_sift_b200.synth import synth_descriptors

    q, t = synth_descriptors(n, n)
    rng = np.random.default_rng(7)
    dq, dt = [], []
    for r in range(rotate):
        pq, pt = rng.permutation(n), rng.permutation(n)
        dq.append(torch.from_numpy(q[pq].copy().view(np.int16)).cuda())
        dt.append(torch.from_numpy(t[pt].copy().view(np.int16)).cuda())
    out = torch.empty(n, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for i in range(warmup):
        matchBruteForce(dq[i % rotate], n, dt[i % rotate], n, stream=s, out=out)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for i, (a, b) in enumerate(ev):
        a.record()
        matchBruteForce(dq[i % rotate], n, dt[i % rotate], n, stream=s, out=out)
        b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)
    # back-to-back throughput figure (no event between calls)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(iters):
        matchBruteForce(dq[i % rotate], n, dt[i % rotate], n, stream=s, out=out)
    b.record()
    torch.cuda.synchronize()
    stream_ms = a.elapsed_time(b) / iters
    # host buffers through the blocking reference-style call (H2D + match + D2H inside)
    for _ in range(3):
        matchBruteForce(q, n, t, n)
    t0 = time.perf_counter()
    reps = 20
    for _ in range(reps):
        res = matchBruteForce(q, n, t, n)
    host_ms = (time.perf_counter() - t0) * 1e3 / reps
    peak, kind = tensor_peak_tflops()
    flops = 2.0 * n * n * 128
    med = ms[len(ms) // 2]
    return {"n": n, "ms_median": med, "ms_min": ms[0], "ms_p90": ms[int(0.9 * len(ms))], "ms_back_to_back": stream_ms,
            "ms_host_buffers_blocking": host_ms, "matches": int((res >= 0).sum()),
            "roofline": {"bound": "tensor", "achieved": flops / (stream_ms * 1e-3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                         "frac": flops / (stream_ms * 1e-3) / 1e12 / peak, "peak_kind": kind}}


def large_sharded(n: int, world: int, rank: int, iters: int = 5, check_rows: int = 64):
    """BASELINE.json configs[4]: n x n descriptors, query rows sharded over the ranks, through the native call
    b200sift_match_sharded. Rank 0 owns the train set: the timed region contains its ncclBroadcast, the matcher on this
    rank's row block and the ncclAllGather of the int32 indices, all on one stream (CUDA events, max over ranks, median
    of `iters`). Correctness: `check_rows` sampled rows of every rank's block against the CPU restatement of
    Match.cu:35-177 (oracle/, bit-exact: the SURVEY 8(c) oracle for this size) + all ranks hold the same result."""
    import time

    import torch

    from another_cuda_sift_b200.sharding import ShardedMatcher
    from another_cuda_sift_b200.synth import synth_descriptors

    q, t = synth_descriptors(n, n)
    sm = ShardedMatcher()
    lo, hi = sm.rows(n)
    dq = torch.from_numpy(q[lo:hi].copy().view(np.int16)).cuda()            # this rank's query rows only
    dt_src = torch.from_numpy(t.view(np.int16)).cuda() if rank == 0 else None  # the train set lives on rank 0
    dt = torch.empty((n, 128), dtype=torch.int16, device="cuda")               # receive buffer (rank 0: working copy)
    root = 0 if world > 1 else -1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
        torch.cuda.synchronize()

    def reset():  # the receivers forget the train set, so that every timed call really needs the broadcast
        if rank == 0:
            dt.copy_(dt_src)
        else:
            dt.zero_()

    times, full = [], None
    for it in range(iters + 2):
        reset()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        full = sm.match(dq, n, dt, n, train_root=root)
        b.record()
        torch.cuda.synchronize()
        ms = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
        if world > 1:
            import torch.distributed as dist

            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if it >= 2:
            times.append(float(ms.item()))
    reset()
    barrier()
    stages = sm.profile(dq, n, dt, n, train_root=root)
    if world > 1:
        import torch.distributed as dist

        st = torch.tensor(list(stages.values()), dtype=torch.float64, device="cuda")
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
        stages = dict(zip(stages.keys(), [float(v) for v in st.tolist()]))
    idx = full.cpu().numpy()
    # correctness against the oracle on sampled rows of THIS rank's block
    from oracle.pyoracle import CpuOracle

    rows = lo + np.sort(np.random.default_rng(100 + rank).choice(hi - lo, size=min(check_rows, hi - lo), replace=False))
    t0 = time.perf_counter()
    cpu = CpuOracle(64, 64, 100)
    if world > 1:
        cpu.set_threads(max(1, (os.cpu_count() or 8) // world))
    want = cpu.match(q[rows], t)
    oracle_s = time.perf_counter() - t0
    bad = int((want != idx[rows]).sum())
    chk = torch.tensor([int(np.bitwise_xor.reduce(idx.astype(np.int64) * 2654435761 % (1 << 31))), bad], dtype=torch.int64, device="cuda")
    same, bad_total = True, bad
    if world > 1:
        import torch.distributed as dist

        lo_c, hi_c = chk[:1].clone(), chk[:1].clone()
        dist.all_reduce(lo_c, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi_c, op=dist.ReduceOp.MAX)
        same = bool(lo_c.item() == hi_c.item())
        tot = chk[1:].clone()
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        bad_total = int(tot.item())
    ms = sorted(times)[len(times) // 2]
    peak, kind = tensor_peak_tflops()
    flops = 2.0 * n * n * 128
    sm.close()
    return {"n": n, "n_gpus": world, "rows_per_rank": hi - lo, "ms": ms, "ms_all": times, "matches": int((idx >= 0).sum()),
            "stages_ms_max_over_ranks": {k: round(v, 4) for k, v in stages.items()},
            "timed_region": "ncclBroadcast of the 25.6 MB train set from rank 0 + matcher on this rank's rows + ncclAllGather of int32 indices, one stream, "
                            "CUDA events, max over ranks" if world > 1 else "matcher on all rows (one rank: no collective)",
            "oracle_check": {"rows_per_rank": int(len(rows)), "ranks": world, "mismatches": bad_total, "oracle": "oracle/sift_oracle.cpp match (Match.cu:35-177 restated, bit-exact)",
                             "oracle_seconds_rank0": round(oracle_s, 1)},
            "all_ranks_identical": same, "broadcast_bytes": n * 256, "gather_bytes_per_rank": 4 * (hi - lo),
            "roofline": {"bound": "tensor", "achieved": flops / (ms * 1e-3) / 1e12, "peak": peak * world, "unit": "TFLOP/s",
                         "frac": flops / (ms * 1e-3) / 1e12 / (peak * world), "peak_kind": kind}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--small", type=int, default=2000)
    ap.add_argument("--large", type=int, default=20000)
    ap.add_argument("--iters", type=int, default=200)
    args = ap.parse_args()
    import torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if os.environ.get("B200SIFT_MATCH_PATH"):  # A/B measurements: 3 = single-pass streaming filter
        from another_cuda_sift_b200 import _lib

        _lib.load().b200sift_match_set_path(int(os.environ["B200SIFT_MATCH_PATH"]))
    line = {"metric": "matchBruteForce latency", "unit": "ms", "n_gpus": world, "data": "synthetic (seed 42/43)", "dtype": "f16 in, f32 accumulate, f64 ratio test"}
    if rank == 0 and args.small > 0:
        line["small"] = match_latency(args.small, args.iters)
    if args.large > 0:
        line["large"] = large_sharded(args.large, world, rank)
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
