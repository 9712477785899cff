"""Multi-GPU partitioning of the SIFT hot path; one process per GPU.

The reference is single-GPU (no NCCL / multi-process code at all), so there is no reference file to
mirror here; the two shardable units of work are

* independent frames of a batch        -> contiguous runs of frames per rank, NO collective on the data
                                          path (each rank keeps its frames' results);
* query rows of a large matching job   -> every rank matches its block of query rows against the whole
                                          train set (the per-query result needs every train column,
                                          quirk Q15), then ONE all-gather of int32 indices.

`torch.distributed` is plumbing only (rendezvous + the NCCL all-gather). The compute is the CUDA
matcher behind the C ABI; tests inject a different `compute` to exercise the plumbing on CPU/gloo.
"""
from __future__ import annotations

from typing import Callable, Tuple


def frame_shard(n_frames: int, world: int, rank: int, overlap: int = 0) -> range:
    """Contiguous run of frames for `rank`. With overlap=1 every run also starts one frame early
    (except the first) so that each consecutive pair (i-1, i) is local to exactly one rank:
    rank r reports matches for pairs ending at its own frames."""
    per, rem = divmod(n_frames, world)
    start = rank * per + min(rank, rem)
    stop = start + per + (1 if rank < rem else 0)
    return range(max(0, start - overlap) if stop > start else start, stop)


def row_shard(n_rows: int, world: int, rank: int) -> Tuple[int, int]:
    """Equal-size blocks of ceil(n/world) query rows (the last ones may be short / empty) so a
    fixed-size all-gather reassembles the result without a size exchange."""
    per = -(-n_rows // world) if n_rows > 0 else 0
    start = min(n_rows, rank * per)
    return start, min(n_rows, start + per)


def sharded_match(des, num_des: int, src, num_src: int, group=None, compute: Callable | None = None):
    """`sift_cuda::matchBruteForce` (reference sift_func/Match.cuh:9-14) with the query rows split over
    the ranks of `group`; every rank returns the full int32[num_des] result.

    `des` / `src` are the full descriptor sets, resident on every rank (torch tensors, float16 or
    int16 view, [n,128]). compute(des_block, n, src, num_src) -> int32 tensor[n] defaults to the CUDA
    matcher on the current stream.
    """
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if compute is None:
        from .detector import matchBruteForce

        def compute(d, n, s, m):
            return matchBruteForce(d, n, s, m, stream=torch.cuda.current_stream().cuda_stream)

    lo, hi = row_shard(num_des, world, rank)
    per = -(-num_des // world) if num_des > 0 else 0
    local = torch.full((max(per, 1),), -1, dtype=torch.int32, device=des.device)
    if hi > lo:
        local[: hi - lo] = compute(des[lo:hi], hi - lo, src, num_src)
    if world == 1:
        return local[:num_des]
    full = torch.empty(world * max(per, 1), dtype=torch.int32, device=des.device)
    dist.all_gather_into_tensor(full, local, group=group)
    return full[:num_des]
