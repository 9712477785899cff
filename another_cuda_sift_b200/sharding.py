"""Multi-GPU partitioning of the SIFT hot path; one process per GPU.

The reference is single-GPU (no NCCL / multi-process code at all), so there is no reference file to
mirror here; the two shardable units of work are

* independent frames of a batch        -> contiguous runs of frames per rank, NO collective on the data
                                          path (each rank keeps its frames' results);
* query rows of a large matching job   -> every rank matches its block of query rows against the whole
                                          train set (the per-query result needs every train column,
                                          quirk Q15): ncclBroadcast of the train set, the matcher,
                                          ncclAllGather of int32 indices, all enqueued on one stream by
                                          the native call b200sift_match_sharded (csrc/shard.cu).

`torch.distributed` is rendezvous only (it ships the 128-byte NCCL id); the communicator, the
collectives and the matcher live behind the C ABI. `sharded_match(..., compute=...)` keeps a pure
torch.distributed path so that the partitioning logic can be tested on CPU with gloo.
"""
from __future__ import annotations

import ctypes as C
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
    fixed-size all-gather reassembles the result without a size exchange (== b200sift_shard_rows)."""
    per = -(-n_rows // world) if n_rows > 0 else 0
    start = min(n_rows, rank * per)
    return start, min(n_rows, start + per)


class ShardedMatcher:
    """`sift_cuda::matchBruteForce` (reference sift_func/Match.cuh:9-14) with the query rows split over the ranks of
    a torch.distributed group, through the native b200sift_match_sharded: the process group is used ONCE, to ship the
    NCCL id of the engine's own communicator; every call then enqueues broadcast + matcher + all-gather on one CUDA
    stream without touching the host again."""

    def __init__(self, group=None):
        import torch
        import torch.distributed as dist

        from . import _lib
        from ._lib import check

        self.lib = _lib.load()
        self._check = check
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        ident = torch.zeros(128, dtype=torch.uint8)
        if self.rank == 0:
            buf = (C.c_uint8 * 128)()
            check(self.lib.b200sift_nccl_unique_id(buf))
            ident = torch.frombuffer(bytearray(buf), dtype=torch.uint8).clone()
        if self.world > 1:
            dev = ident.cuda() if dist.get_backend(group) == "nccl" else ident
            dist.broadcast(dev, src=0, group=group)
            ident = dev.cpu()
        raw = (C.c_uint8 * 128).from_buffer_copy(bytes(ident.numpy().tobytes()))
        h = C.c_void_p()
        check(self.lib.b200sift_comm_create(raw, self.rank, self.world, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self.lib.b200sift_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def rows(self, n_rows: int) -> Tuple[int, int]:
        return row_shard(n_rows, self.world, self.rank)

    def match(self, des_block, num_des_total: int, src, num_src: int, train_root: int = -1, out=None, stream: int | None = None):
        """des_block: this rank's rows [lo, hi) of the query set (torch CUDA tensor, float16 / int16 view, [hi-lo, 128]);
        src: the train set buffer [num_src, 128] on this rank -- its contents only matter on `train_root` (the other
        ranks receive them; train_root = -1: every rank already holds the set). Returns an int32 CUDA tensor with the
        full result (num_des_total entries) on every rank, enqueued on `stream` (default: torch's current stream)."""
        import torch

        per = -(-num_des_total // self.world) if num_des_total > 0 else 0
        if out is None:
            out = torch.empty(max(self.world * per, 1), dtype=torch.int32, device=src.device)
        if stream is None:
            stream = torch.cuda.current_stream().cuda_stream
        self._check(self.lib.b200sift_match_sharded(self._h, des_block.data_ptr() if des_block is not None and des_block.numel() else None,
                                                    num_des_total, src.data_ptr(), num_src, train_root, out.data_ptr(), C.c_void_p(stream)))
        return out[:num_des_total]

    def profile(self, des_block, num_des_total: int, src, num_src: int, train_root: int = -1) -> dict:
        """One blocking call with an event between the stages: ms per stage on this rank."""
        import torch

        per = -(-num_des_total // self.world) if num_des_total > 0 else 0
        out = torch.empty(max(self.world * per, 1), dtype=torch.int32, device=src.device)
        ms = (C.c_float * 5)()
        self._check(self.lib.b200sift_match_sharded_profile(
            self._h, des_block.data_ptr() if des_block is not None and des_block.numel() else None, num_des_total, src.data_ptr(), num_src,
            train_root, out.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream), ms))
        return dict(zip(("broadcast", "norms", "filter", "rerank", "allgather"), [float(v) for v in ms]))


def sharded_match(des, num_des: int, src, num_src: int, group=None, compute: Callable | None = None):
    """Query-row sharded matchBruteForce with the plumbing in torch.distributed: every rank matches its block with
    `compute(des_block, n, src, num_src) -> int32 tensor[n]` and one all-gather reassembles the result. This is the
    CPU-testable statement of the partitioning (tests/test_host_logic.py runs it on gloo with an injected compute); on
    GPUs use ShardedMatcher, which does the same with native NCCL calls on the matcher's stream.

    `des` / `src` are the full descriptor sets, resident on every rank (torch tensors, float16 or
    int16 view, [n,128])."""
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
