#!/usr/bin/env python
"""bench.py -- detectAndCompute frames/s @1920x1200 (+ consecutive-frame matching), BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one batch of --frames 1920x1200 synthetic frames through the hot path on every rank:
detectAndCompute on each frame, then sift_cuda::matchBruteForce(previous frame, this frame) exactly as
the reference's tool/extract_and_match_example.cc:57-101 drives it. Frames of a batch are independent units:
each rank has its own frames and there is NO collective on the data path ("scaling": "weak").

  value  frames/s over all ranks, inputs resident in HBM (device-resident frames > L2 per step)
  e2e    the same through the reference-facing API with HOST buffers: detectAndCompute(host image) ->
         results (keypoints, features, descriptors, match indices) to pinned host memory, through the batched
         API Detector.detectAndComputeBatch (H2D of frame i+1 overlaps the graph of frame i), copies inside the timing
  roofline       dominant kernel class, algorithmic bytes / CUDA-event time vs MEASURED_PEAKS.json
  cpu_baseline   the CPU restatement (oracle/) timed on this box's host cores on a bounded sample
  --impl reference   the UNMODIFIED reference CUDA build (oracle/_ref, its own public API and kernels)
                     on the same config; the reference has no CPU implementation of this path, so when
                     oracle/_ref is absent the CPU oracle port is timed instead ("Measurement").

Exactly one JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "detectAndCompute frames/s @1920x1200"
AFFINITY = "off"
COLS, ROWS = 1920, 1200
NUM_FEATURES = 5000
S_PIXELS = 3072033  # sum of pixels over the 10 octaves of 1920x1200


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), float(j.get("bf16_tflops_sustained", j.get("bf16_tflops", 1400.0))), "measured"
    return 6650.0, 1400.0, "fallback"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons of one GPU sampled during the timed region (NVML)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            log("clock sampler unavailable:", e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            self._stop_evt.wait(0.05)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def make_frames(n: int, rank: int):
    """A run of `n` consecutive synthetic frames: textured base frame, then each frame
    is the previous one shifted 2 px / rotated 0.5 deg + noise, so consecutive-frame matching is real."""
    from another_cuda_sift_b200.synth import next_frame, synth_frame

    frames = [synth_frame(ROWS, COLS, 1000 * rank)]
    for i in range(1, n):
        frames.append(next_frame(frames[-1], 1000 * rank + i))
    return frames


def bind_to_gpu_numa(local: int) -> str:
    """One process per GPU: pin this process (and the threads CUDA / torch start later) to the CPU cores NVML
    reports as local to the GPU, so that the pinned host frames and result buffers are allocated on the GPU's own
    NUMA node and its DMA traffic does not cross the socket interconnect. Without it the end-to-end (host buffer)
    figure stops scaling at 8 GPUs while the device-resident one scales 7.9x."""
    if os.environ.get("B200SIFT_NO_AFFINITY"):
        return "off"
    try:
        import pynvml

        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
        return f"{len(os.sched_getaffinity(0))} cores"
    except Exception as e:  # noqa: BLE001
        return f"unavailable ({e})"


def dist_setup(gpus: int):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    global AFFINITY
    AFFINITY = bind_to_gpu_numa(local)  # before the CUDA context and every pinned allocation
    import torch

    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(0)
    if gpus != world:
        log(f"note: --gpus {gpus} but WORLD_SIZE={world}; using WORLD_SIZE")
    return world, rank, local


def barrier_sync(world):
    import torch

    torch.cuda.synchronize()
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(ms: float, world: int) -> float:
    if world == 1:
        return ms
    import torch
    import torch.distributed as dist

    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(v: float, world: int) -> float:
    if world == 1:
        return v
    import torch
    import torch.distributed as dist

    t = torch.tensor([v], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def run_b200(args, world, rank, local):
    """Our arm: device-resident `value`, host-buffer `e2e`, roofline of the dominant kernel class."""
    import torch

    from another_cuda_sift_b200 import CudaSiftConfig, Detector, matchBruteForce
    from another_cuda_sift_b200 import _lib

    lib = _lib.load()
    F = args.frames
    frames = make_frames(F, rank)
    det = Detector(CudaSiftConfig(col_width=COLS, row_width=ROWS, numFeatures=NUM_FEATURES))
    assert det.gpuWarmUpAndAllocate()
    stream = torch.cuda.ExternalStream(det.stream)
    d_frames = [torch.from_numpy(f).cuda() for f in frames]  # F x 9.2 MB resident in HBM (> L2 for F >= 14)
    h_frames = [torch.from_numpy(f).pin_memory() for f in frames]
    idx_dev = [torch.empty(NUM_FEATURES, dtype=torch.int32, device="cuda") for _ in range(det.frames_in_flight)]
    idx_host = torch.empty(NUM_FEATURES, dtype=torch.int32).pin_memory()
    launches_frame = det.launches_per_frame()
    state = {"prev": 0, "kpts": 0, "launches": 0, "d2h": 0}

    from another_cuda_sift_b200 import HostResults

    results = [HostResults(NUM_FEATURES) for _ in range(F)]
    match_host = [torch.empty(NUM_FEATURES, dtype=torch.int32).pin_memory() for _ in range(F)]
    h_np = [h.numpy() for h in h_frames]
    n_match_launches = lib.b200sift_match_launches(NUM_FEATURES, NUM_FEATURES)

    def step_resident():
        # inputs already in HBM; frame graph + consecutive-frame matcher per frame, sizes stay on the device,
        # one host synchronisation per step
        for i in range(F):
            det.detectAndComputeAsync(d_frames[i])
            det.matchConsecutiveAsync(idx_dev[i % len(idx_dev)])
        det.synchronize()
        state["launches"] += F * (launches_frame + n_match_launches)

    def step_e2e():
        # the batched-frame public API with HOST buffers: pinned H2D of every frame (overlapping the previous
        # frame's graph), keypoints + features + descriptors + match indices back to pinned host memory
        det.detectAndComputeBatch(h_np, results, match_host)
        state["d2h"] += F * (results[0].nbytes() + NUM_FEATURES * 4)

    # the same call fed with the 8-bit frames the reference's float images are cast from (cvUtils/ConversionImpl.hpp:25-27;
    # the synthetic frames are integer valued, results are bit-identical: tests/test_gpu_parity.py): 4x fewer H2D bytes
    h8_frames = [torch.from_numpy(f.astype(np.uint8)).pin_memory() for f in frames]
    h8_np = [h.numpy() for h in h8_frames]

    def step_e2e_u8():
        det.detectAndComputeBatch(h8_np, results, match_host)

    def timed(step_fn, steps, warmup):
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                step_fn()
            barrier_sync(world)
            state.update(kpts=0, launches=0, d2h=0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record(stream)
            for _ in range(steps):
                step_fn()
            e1.record(stream)
            barrier_sync(world)
            wall = (time.perf_counter() - t0) * 1e3
            ms = e0.elapsed_time(e1)
        return max_over_ranks(ms, world), wall

    sampler = ClockSampler(local)
    sampler.start()
    ms_total, wall = timed(step_resident, args.steps, args.warmup)
    clocks = sampler.stop()
    kpts_per_frame = 0.0
    launches = state["launches"]
    ms_e2e, _ = (ms_total, 0) if args.quick else timed(step_e2e, args.steps, args.warmup)
    d2h_per_step = state["d2h"] / args.steps if not args.quick else 0
    ms_e2e_u8, _ = (ms_total, 0) if args.quick else timed(step_e2e_u8, args.steps, args.warmup)
    if not args.quick:
        kpts_per_frame = float(np.mean([r.total for r in results]))

    total_frames = sum_over_ranks(F * args.steps, world)
    value = total_frames / (ms_total * 1e-3)
    e2e_value = total_frames / (ms_e2e * 1e-3)
    e2e_u8_value = total_frames / (ms_e2e_u8 * 1e-3)

    # pure frame-graph time (no host sync per frame, no matcher) and per-class breakdown, rank 0 only
    extra = {}
    if rank == 0:
        det.detectAndCompute(d_frames[0])
        graph_ms = det.time_frames(5, 5 if args.quick else 100)
        pipe_ms = det.time_frames_pipelined(6, 6 if args.quick else 120)
        prof = det.profile_frame(1 if args.quick else 3, 1 if args.quick else 20)
        extra = dict(graph_ms=graph_ms, pipe_ms=pipe_ms, prof=prof, kpts=det.total_size, stages=det.time_stages(5 if args.quick else 50))
    return dict(value=value, ms_per_step=ms_total / args.steps, wall_ms=wall, e2e_value=e2e_value, e2e_u8_value=e2e_u8_value,
                e2e_ms_per_step=ms_e2e / args.steps, clocks=clocks, kpts_per_frame=kpts_per_frame,
                launches=launches, d2h_per_step=d2h_per_step, frames=frames, det=det, extra=extra)


def cpu_port_baseline(frames, budget_s: float = 12.0):
    """The CPU restatement of the reference path (oracle/sift_oracle.cpp: scalar C++ spread over all host
    threads) on a bounded sample of the same workload: detect on consecutive 1920x1200 frames + match."""
    from oracle.pyoracle import CpuOracle

    cpu = CpuOracle(COLS, ROWS, NUM_FEATURES)
    cores = cpu.threads()
    n_done, t0, prev = 0, time.perf_counter(), None
    while n_done < len(frames) and (time.perf_counter() - t0 < budget_s or n_done < 2):
        n = cpu.detect(frames[n_done])
        cur = cpu.results()["desc"][:n].copy()
        if prev is not None and n > 0 and len(prev) > 0:
            cpu.match(prev, cur)
        prev = cur
        n_done += 1
    dt = time.perf_counter() - t0
    return dict(value=n_done / dt, unit="frames/s", cores=cores, kind="port",
                sample=f"{n_done} consecutive {COLS}x{ROWS} frames (detect + match vs previous), {dt:.1f} s, oracle/sift_oracle.cpp")


def opencv_baseline(frames, budget_s: float = 10.0):
    """cv::SIFT(5000,3,0.04,10,1.6) + BFMatcher(L2) knn-2 + 0.8 ratio on the host cores (reported only)."""
    try:
        import cv2
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"cv2 import failed: {e}"}
    sift = cv2.SIFT_create(5000, 3, 0.04, 10, 1.6)
    bf = cv2.BFMatcher(cv2.NORM_L2)
    n_done, t0, prev = 0, time.perf_counter(), None
    while n_done < len(frames) and (time.perf_counter() - t0 < budget_s or n_done < 2):
        _, des = sift.detectAndCompute(frames[n_done].astype(np.uint8), None)
        if prev is not None and des is not None and len(des) > 1:
            [m for m in bf.knnMatch(prev, des, k=2) if len(m) == 2 and m[0].distance < 0.8 * m[1].distance]
        prev = des
        n_done += 1
    dt = time.perf_counter() - t0
    return dict(value=n_done / dt, unit="frames/s", cores=cv2.getNumThreads(), version=cv2.__version__,
                sample=f"{n_done} frames, cv2.SIFT_create(5000,3,0.04,10,1.6) (2x upscale on) + BFMatcher knn2 + ratio 0.8")


def run_reference(args, world, rank, local):
    """The reference arm. The reference implements this path ONLY in CUDA (no CPU code), so its 'own
    implementation' is its CUDA build: oracle/_ref/libref_sift.so, compiled from the untouched sources for
    sm_100a, driven through Detector::detectAndCompute / copyToHost / matchBruteForce. Rank 0 only."""
    from oracle.pyoracle import RefOracle, ref_available
    import ctypes as C

    F = args.frames
    frames = make_frames(F, 0)
    if not ref_available():
        b = cpu_port_baseline(frames, budget_s=20.0)
        ms_step = F / b["value"] * 1e3
        return dict(value=b["value"], ms_per_step=ms_step, e2e_value=b["value"], cpu_baseline=b, frames_per_step=F,
                    note="oracle/_ref missing: CPU oracle port timed instead", clocks=None)
    ref = RefOracle(COLS, ROWS, NUM_FEATURES)
    lib = ref.lib
    lib.ref_add_frame.restype, lib.ref_add_frame.argtypes = C.c_int, [C.c_void_p, C.c_void_p]
    lib.ref_process_frame.restype, lib.ref_process_frame.argtypes = C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    lib.ref_process_resident.restype, lib.ref_process_resident.argtypes = C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    lib.ref_upload_frame.restype, lib.ref_upload_frame.argtypes = None, [C.c_void_p, C.c_int]
    for f in frames:
        lib.ref_add_frame(ref.h, np.ascontiguousarray(f).ctypes.data)
    nm = C.c_int(0)
    sampler = ClockSampler(local)

    def step_e2e():
        for i in range(F):
            lib.ref_process_frame(ref.h, i, 1, 1, C.byref(nm))

    for _ in range(args.warmup):
        step_e2e()
    sampler.start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    ms_e2e = (time.perf_counter() - t0) * 1e3  # every call ends in cudaDeviceSynchronize: wall == device
    clocks = sampler.stop()
    # device-resident: the reference cannot take a device image, so one frame is uploaded outside the
    # timing and the stage sequence of detectAndCompute is replayed on it (no H2D), + matcher.
    lib.ref_upload_frame(ref.h, 0)
    for _ in range(args.warmup * 4):
        lib.ref_process_resident(ref.h, 1, C.byref(nm))
    t0 = time.perf_counter()
    for _ in range(args.steps * F):
        lib.ref_process_resident(ref.h, 1, C.byref(nm))
    ms_res = (time.perf_counter() - t0) * 1e3
    total = F * args.steps
    return dict(value=total / (ms_res * 1e-3), ms_per_step=ms_res / args.steps, e2e_value=total / (ms_e2e * 1e-3),
                frames_per_step=F, clocks=clocks, kpts=int(lib.ref_total_size(ref.h)),
                cpu_baseline=dict(value=total / (ms_e2e * 1e-3), unit="frames/s", cores=1, kind="reference",
                                  sample=f"{total} frames; the reference has no CPU implementation of this path: this is its "
                                         "unmodified CUDA build (oracle/_ref, sm_100a) on the same B200, 1 host thread driving it"))


def class_bytes(kpts: int):
    """Algorithmic HBM bytes per frame of each kernel class (DESIGN.md 'Kernels'): every blur launch reads
    one layer and writes one layer (8 B/px); detect reads the 6 Gaussian layers once (24 B/px) and writes 3
    bit planes; describe writes 284 B per keypoint (B = 4WH + 48S + 284N per frame)."""
    wh = COLS * ROWS
    return {
        "blur_oct0": 6 * 8 * wh,
        "blur_small": 5 * 8 * (S_PIXELS - wh),
        "downsample": 5 * (S_PIXELS - wh),  # reads 4 source px (16 B) per output px... of the previous octave: 4 B/px in + 4 B/px out of the new octave ~ 5 B per px of octaves >= 1
        "detect": 24 * S_PIXELS + 3 * S_PIXELS // 8,
        "describe": 284 * kpts,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--frames", type=int, default=16, help="frames per step per GPU")
    ap.add_argument("--size", default="1920x1200", help="frame size WxH: 1920x1200 is BASELINE.json's metric; 752x480 and 1600x900 are its other configs")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--quick", action="store_true", help="profiler runs: skip e2e leg, per-class profile and CPU baselines")
    args = ap.parse_args()
    global COLS, ROWS, S_PIXELS, METRIC
    COLS, ROWS = (int(v) for v in args.size.lower().split("x"))
    if (COLS, ROWS) != (1920, 1200):
        METRIC = f"detectAndCompute frames/s @{COLS}x{ROWS}"
        n_oct = int(round(np.log(2.0 * min(COLS, ROWS)) / np.log(2.0) - 2) + 1)          # Detector.hh:27
        rnd = lambda v: int(np.floor(v + 0.5))                                            # roundf, Detector.cu:83
        S_PIXELS = sum(rnd(COLS / 2 ** o) * rnd(ROWS / 2 ** o) for o in range(n_oct))
    args.warmup = max(args.warmup, 3)

    if args.impl == "reference":
        rank = int(os.environ.get("RANK", "0"))
        if rank != 0:
            return 0
        r = run_reference(args, 1, 0, 0)
        line = {
            "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "frames/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{COLS}x{ROWS} detectAndCompute + consecutive-frame matchBruteForce", "frames_per_step_per_gpu": args.frames,
                       "numFeatures": NUM_FEATURES, "reference": "unmodified reference CUDA build (oracle/_ref) on 1 B200; "
                       "the reference is single-GPU, so the same 1-GPU figure is reported for every N"},
            "e2e": {"value": r["e2e_value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "cpu_baseline": r["cpu_baseline"], "clocks": r.get("clocks"),
        }
        if "note" in r:
            line["note"] = r["note"]
        print(json.dumps(line), flush=True)
        return 0

    world, rank, local = dist_setup(args.gpus)
    r = run_b200(args, world, rank, local)
    launches_total = sum_over_ranks(r["launches"], world)
    if rank != 0:
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
            dist.destroy_process_group()
        return 0
    hbm_peak, _, peak_kind = measured_peaks()
    ex = r["extra"]
    prof = ex["prof"]
    cb = class_bytes(ex["kpts"])
    # Dominant kernel FUNCTION of the frame by device time: blur_vh_kernel (the 6 octave-0 launches and the 21
    # launches of the smaller octaves are two profiling classes of the same kernel). Bytes and time are the
    # totals over its launches of one frame, i.e. bytes per launch / average launch duration.
    groups = {"blur_vh_kernel": ["blur_oct0", "blur_small"], "detect_kernel": ["detect"], "descriptor_kernel": ["describe"]}
    gms = {g: sum(prof[k]["ms"] for k in ks) for g, ks in groups.items()}
    dom = max(gms, key=gms.get)
    dom_ms = gms[dom]
    eager_ms = dom_ms
    how = ("CUDA events around every launch of one frame run eagerly on a single stream (b200sift_profile_frame); "
           "in the frame graph these launches overlap")
    if dom == "blur_vh_kernel":
        # The blur launches as the product runs them: the frame GRAPH truncated after the pyramid, replayed and timed with
        # CUDA events on its stream (b200sift_time_stages). That interval also holds the 5 downsample launches and the tail
        # kernel on the blur chain's critical path, so the blur rate derived from it is a lower bound; the eager per-launch
        # figure (5-6 us of launch latency per dependent small kernel) is kept as eager_ms_per_frame.
        dom_ms = ex["stages"]["pyramid"]
        how = ("the frame graph truncated after the pyramid (27 blur launches + 5 downsample + tail kernel), replayed and timed "
               "with CUDA events on its stream (b200sift_time_stages): a lower bound of the blur rate; eager_ms_per_frame is the "
               "sum over the same launches run eagerly with an event after each")
    dom_bytes = sum(cb[k] for k in groups[dom])
    dom_launches = sum(prof[k]["launches"] for k in groups[dom])
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get(dom)
    frame_bytes = 4 * COLS * ROWS + 48 * S_PIXELS + 284 * ex["kpts"]
    line = {
        "metric": METRIC, "value": r["value"], "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{COLS}x{ROWS} detectAndCompute + consecutive-frame matchBruteForce", "frames_per_step_per_gpu": args.frames,
                   "global_frames_per_step": args.frames * world, "numFeatures": NUM_FEATURES, "parallelism": f"frames x{world}", "cpu_affinity": AFFINITY, "frames_in_flight_per_gpu": r["det"].frames_in_flight,
                   "keypoints_per_frame": round(r["kpts_per_frame"], 1),
                   "l2": f"device-resident inputs are {args.frames} distinct frames = {args.frames * COLS * ROWS * 4 / 1e6:.0f} MB per step; with the six pyramids in flight "
                         f"({6 * 24 * S_PIXELS / 1e6:.0f} MB) the working set is {'larger' if (args.frames * COLS * ROWS * 4 + 6 * 24 * S_PIXELS) > 126e6 else 'SMALLER'} than the 126 MB L2"},
        "ms_per_frame": r["ms_per_step"] / args.frames,
        "frame_graph_ms": ex["graph_ms"],             # latency of one frame's graphs, nothing else in flight
        "frame_pipelined_ms": ex["pipe_ms"],          # per frame with every frame slot in flight (graphs only)
        "e2e": {"value": r["e2e_value"], "unit": "frames/s", "h2d_bytes_per_step": args.frames * COLS * ROWS * 4,
                "d2h_bytes_per_step": int(r["d2h_per_step"]), "ms_per_frame": r["e2e_ms_per_step"] / args.frames},
        "e2e_u8": {"value": r["e2e_u8_value"], "unit": "frames/s", "h2d_bytes_per_step": args.frames * COLS * ROWS,
                   "note": "same public call with the 8-bit host frames the float images are cast from (b200sift_detect_host_u8_async): "
                           "identical results, a quarter of the H2D bytes; the float figure above is the headline"},
        "gpu_launches": int(launches_total),
        "clocks": r["clocks"],
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "peak_kind": peak_kind,
                     "launches_per_frame": dom_launches, "ms_per_frame": dom_ms,
                     "bytes_per_launch": dom_bytes / max(dom_launches, 1), "us_per_launch": dom_ms * 1e3 / max(dom_launches, 1),
                     "by_kernel": {g: {"ms_per_frame": round(gms[g], 5),
                                       "GB/s": round(sum(cb[k] for k in groups[g]) / (gms[g] * 1e-3) / 1e9, 1)} for g in groups},
                     "eager_ms_per_frame": eager_ms, "note": how},
        "roofline_frame": {"bytes": frame_bytes, "achieved": frame_bytes / (ex["pipe_ms"] * 1e-3) / 1e9,
                           "frac": frame_bytes / (ex["pipe_ms"] * 1e-3) / 1e9 / hbm_peak,
                           "note": "SURVEY 8(d) algorithmic bytes of one frame / pipelined frame time"},
        "kernel_ms_per_frame": {k: round(v["ms"], 5) for k, v in prof.items()},
        "graph_cumulative_ms": {k: round(v, 5) for k, v in ex["stages"].items()},
    }
    if not args.quick:
        # second half of BASELINE.json's metric: latency of the 2000 x 2000 x 128 ratio-test match (seeded synthetic
        # descriptor sets, device-resident, median of per-call CUDA-event times) + its tensor roofline
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from bench_match import match_latency

        line["match_2k2k"] = match_latency(2000, iters=200)
    if world == 1 and not args.no_cpu_baseline and not args.quick:
        line["cpu_baseline"] = cpu_port_baseline(r["frames"])
        line["opencv_cpu"] = opencv_baseline(r["frames"])
    print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
