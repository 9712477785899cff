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

# The engine's own default (another_cuda_sift_b200/__init__.py, csrc/engine.cu: b200sift_on_load), stated here because it
# must be in the environment before the first CUDA call of the process, and so that BOTH arms run under the same setting:
# 32 hardware work queues instead of CUDA's default 8 (profiles/r2_timeline.md). A value from the caller's environment wins.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

METRIC = "detectAndCompute frames/s @1920x1200"
AFFINITY = "off"
COLS, ROWS = 1920, 1200
NUM_FEATURES = 5000
S_PIXELS = 3072033  # sum of pixels over the 10 octaves of 1920x1200


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# The contract is exactly ONE JSON line on stdout. Libraries print there too (NCCL's version banner, the reference's
# constructor): file descriptor 1 is pointed at stderr for the whole run and the line is written to the saved descriptor.
_REAL_STDOUT = None


def protect_stdout():
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


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


DISTINCT_FRAMES = 16  # distinct frames of a step's sequence (round 1's workload: 147 MB at 1920x1200, ~3800 keypoints per frame)


def make_frames(n: int, rank: int, rows: int = None, cols: int = None):
    """A run of `n` consecutive synthetic frames: textured base frame, then each frame
    is the previous one shifted 2 px / rotated 0.5 deg + noise, so consecutive-frame matching is real."""
    from another_cuda_sift_b200.synth import next_frame, synth_frame

    rows, cols = rows or ROWS, cols or COLS
    frames = [synth_frame(rows, cols, 1000 * rank)]
    for i in range(1, n):
        frames.append(next_frame(frames[-1], 1000 * rank + i))
    return frames


def frame_order(F: int, n: int = DISTINCT_FRAMES):
    """Indices of the F frames of one step into the n distinct frames: the sequence walked forwards and backwards
    (0, 1, .., n-1, n-2, .., 1, 0, 1, ..), so that every pair of consecutive frames is a real consecutive pair and a longer
    step does NOT change the workload. (Every bilinear resample of next_frame() blurs the sequence a little: a step of 64
    distinct frames averages 2600 keypoints per frame at 1920x1200, one of 128 frames 1960, against 3780 for the 16 frames
    round 1 measured on -- a longer run of distinct frames is an easier workload, not a better engine.)"""
    n = min(n, F)
    if n <= 1:
        return [0] * F
    period = list(range(n)) + list(range(n - 2, 0, -1))
    return [period[i % len(period)] for i in range(F)]


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


def measure_size(cols, rows, F, steps, warmup, world, rank, local, clocks=False, quick=False):
    """One workload (frame size) on this rank's GPU: `steps` timed steps of F frames each, three ways --
    resident (device frames, detectAndComputeAsync + matchConsecutiveAsync per frame, one sync per step),
    e2e (detectAndComputeBatch on pinned HOST float frames, all results back to pinned host memory) and e2e_u8 (the
    same call with 8-bit frames). Every step is bracketed by its own CUDA events (median / min / max are reported
    next to the total the throughput is computed from); barrier + synchronize around the timed region, max over ranks."""
    import torch

    from another_cuda_sift_b200 import CudaSiftConfig, Detector, HostResults
    from another_cuda_sift_b200 import _lib

    lib = _lib.load()
    frames = make_frames(min(F, DISTINCT_FRAMES), rank, rows, cols)
    order = frame_order(F)
    slots = int(os.environ.get("B200SIFT_BENCH_SLOTS", "0"))  # development switch: frames in flight (0: the engine's default)
    det = Detector(CudaSiftConfig(col_width=cols, row_width=rows, numFeatures=NUM_FEATURES), **({"frames_in_flight": slots} if slots else {}))
    assert det.gpuWarmUpAndAllocate()
    stream = torch.cuda.ExternalStream(det.stream)
    d_distinct = [torch.from_numpy(f).cuda() for f in frames]
    h_distinct = [torch.from_numpy(f).pin_memory().numpy() for f in frames]
    h8_distinct = [torch.from_numpy(f.astype(np.uint8)).pin_memory().numpy() for f in frames]
    d_frames = [d_distinct[k] for k in order]   # the F frames of a step: every one is copied / processed again
    h_np = [h_distinct[k] for k in order]
    h8_np = [h8_distinct[k] for k in order]
    idx_dev = [torch.empty(NUM_FEATURES, dtype=torch.int32, device="cuda") for _ in range(det.frames_in_flight)]
    results = [HostResults(NUM_FEATURES) for _ in range(F)]
    match_host = [torch.empty(NUM_FEATURES, dtype=torch.int32).pin_memory() for _ in range(F)]
    launches_frame = det.launches_per_frame()
    n_match_launches = lib.b200sift_match_launches(NUM_FEATURES, NUM_FEATURES)
    state = {"launches": 0}

    def step_resident():
        for i in range(F):
            det.detectAndComputeAsync(d_frames[i])
            det.matchConsecutiveAsync(idx_dev[i % len(idx_dev)])
        det.synchronize()
        state["launches"] += F * (launches_frame + n_match_launches)

    def step_resident_nomatch():
        for i in range(F):
            det.detectAndComputeAsync(d_frames[i])
        det.synchronize()

    def step_e2e():
        det.detectAndComputeBatch(h_np, results, match_host)

    def step_e2e_u8():
        det.detectAndComputeBatch(h8_np, results, match_host)

    def timed(step_fn, steps, warmup):
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                step_fn()
            barrier_sync(world)
            state["launches"] = 0
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
            ev[0].record(stream)
            for k in range(steps):
                step_fn()
                ev[k + 1].record(stream)
            barrier_sync(world)
            ms = ev[0].elapsed_time(ev[steps])
            per = sorted(ev[k].elapsed_time(ev[k + 1]) for k in range(steps))
        return max_over_ranks(ms, world), per

    sampler = ClockSampler(local) if clocks else None
    if sampler:
        sampler.start()
    ms_res, per_res = timed(step_resident, steps, warmup)
    clk = sampler.stop() if sampler else None
    launches = state["launches"]
    out = dict(cols=cols, rows=rows, frames_per_step=F, steps=steps, det=det, frames=frames, clocks=clk, launches=launches,
               ms_res=ms_res, per_res=per_res)
    if quick:
        out.update(ms_e2e=ms_res, ms_u8=ms_res, ms_nomatch=ms_res, kpts=0.0, d2h=0)
        return out
    ms_e2e, per_e2e = timed(step_e2e, steps, warmup)
    kpts = float(np.mean([r.total for r in results]))
    ms_u8, _ = timed(step_e2e_u8, steps, warmup)
    ms_nomatch, _ = timed(step_resident_nomatch, steps, warmup)
    out.update(ms_e2e=ms_e2e, per_e2e=per_e2e, ms_u8=ms_u8, ms_nomatch=ms_nomatch, kpts=kpts,
               d2h=F * (results[0].nbytes() + NUM_FEATURES * 4))
    return out


def size_summary(m, world):
    """The per-configuration block of the JSON line (all ranks' frames over the max-over-ranks time)."""
    F, steps = m["frames_per_step"], m["steps"]
    total = sum_over_ranks(F * steps, world)
    per_frame = lambda ms: ms / steps / F
    med = lambda v: v[len(v) // 2]
    d = {"frames_per_step_per_gpu": F, "steps": steps,
         "value": total / (m["ms_res"] * 1e-3), "ms_per_frame": per_frame(m["ms_res"]),
         "ms_per_step_median": med(m["per_res"]), "ms_per_step_min": m["per_res"][0], "ms_per_step_max": m["per_res"][-1],
         "e2e": total / (m["ms_e2e"] * 1e-3), "e2e_ms_per_frame": per_frame(m["ms_e2e"]),
         "e2e_u8": total / (m["ms_u8"] * 1e-3),
         "keypoints_per_frame": round(m["kpts"], 1),
         "h2d_bytes_per_step": F * m["cols"] * m["rows"] * 4, "d2h_bytes_per_step": int(m["d2h"]),
         # where the time of a frame goes with every slot in flight (exposed cost of each part, this rank):
         "breakdown_ms_per_frame": {"detect": per_frame(m["ms_nomatch"]), "match": per_frame(m["ms_res"] - m["ms_nomatch"]),
                                    "host_copies_not_hidden": per_frame(m["ms_e2e"] - m["ms_res"])}}
    return d


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


def opencv_config0(budget_s: float = 8.0):
    """BASELINE.json configs[0]: 752x480 single-frame cv::SIFT detectAndCompute + a 2000 x 2000 128-D ratio-test match
    via cv::BFMatcher (knn-2, 0.8), on this box's host cores (thirdparty/opencv/modules/features2d/src/sift.dispatch.cpp,
    matchers.cpp:757 in the reference tree; here the cv2 wheel of the image). Reported, not optimised."""
    try:
        import cv2
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"cv2 import failed: {e}"}
    from another_cuda_sift_b200.synth import synth_descriptors

    frames = make_frames(4, 0, 480, 752)
    sift = cv2.SIFT_create(5000, 3, 0.04, 10, 1.6)
    ms, n_kp, t_start = [], 0, time.perf_counter()
    for i in range(12):
        t0 = time.perf_counter()
        kp, _ = sift.detectAndCompute(frames[i % 4].astype(np.uint8), None)
        ms.append((time.perf_counter() - t0) * 1e3)
        n_kp = len(kp)
        if time.perf_counter() - t_start > budget_s and i >= 3:
            break
    q, t = synth_descriptors(2000, 2000)
    q32, t32 = q.astype(np.float32), t.astype(np.float32)
    bf = cv2.BFMatcher(cv2.NORM_L2)
    mm, good = [], 0
    for _ in range(5):
        t0 = time.perf_counter()
        pairs = bf.knnMatch(q32, t32, k=2)
        good = sum(1 for m in pairs if len(m) == 2 and m[0].distance < 0.8 * m[1].distance)
        mm.append((time.perf_counter() - t0) * 1e3)
    return {"workload": "752x480 cv2.SIFT_create(5000,3,0.04,10,1.6).detectAndCompute (2x upscale on) + BFMatcher(NORM_L2).knnMatch(k=2) "
                        "2000 x 2000 x 128 + 0.8 ratio", "cores": cv2.getNumThreads(), "version": cv2.__version__,
            "detect_ms_median": sorted(ms)[len(ms) // 2], "detect_frames_per_s": 1e3 / (sum(ms) / len(ms)), "keypoints": n_kp,
            "match_2k2k_ms_median": sorted(mm)[len(mm) // 2], "match_accepted": good}


def workload_config(cols, rows, frames, world):
    """The `config` keys both arms print (identical for the same workload, so the driver can compare them)."""
    return {"workload": f"{cols}x{rows} detectAndCompute + consecutive-frame matchBruteForce", "frames_per_step_per_gpu": frames,
            "distinct_frames": min(frames, DISTINCT_FRAMES),
            "frame_order": "the distinct frames walked forwards and backwards (every consecutive pair is a real consecutive pair; a longer step is the same workload)",
            "numFeatures": NUM_FEATURES, "numOctaveLayers": 3, "sigma": 1.6, "data": "synthetic frames, seed 1234 + 1000*rank + i",
            "cuda_device_max_connections": os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS")}


def run_reference(args, world, rank, local):
    """The reference arm. The reference implements this path ONLY in CUDA (no CPU code), so its 'own
    implementation' is its CUDA build: oracle/_ref/libref_sift.so, compiled from the untouched sources for
    sm_100a. e2e = its stock public path (Detector::detectAndCompute(Imagef) -> copyToHost(true) -> matchBruteForce,
    tool/extract_and_match_example.cc:57-101) over the SAME distinct frames as our arm; value = the same frames with the
    host->device upload and copyToHost outside the timing (stage sequence of detectAndCompute on the resident image +
    matchBruteForce). Rank 0 only."""
    from oracle.pyoracle import RefOracle, ref_available
    import ctypes as C

    F = args.frames
    frames = make_frames(min(F, DISTINCT_FRAMES), 0)
    order = frame_order(F)
    if not ref_available():
        b = cpu_port_baseline(frames, budget_s=20.0)
        ms_step = F / b["value"] * 1e3
        return dict(value=b["value"], ms_per_step=ms_step, e2e_value=b["value"], cpu_baseline=b, frames_per_step=F,
                    note="oracle/_ref missing: CPU oracle port timed instead", clocks=None)
    ref = RefOracle(COLS, ROWS, NUM_FEATURES)
    lib = ref.lib
    lib.ref_add_frame.restype, lib.ref_add_frame.argtypes = C.c_int, [C.c_void_p, C.c_void_p]
    lib.ref_process_frame.restype, lib.ref_process_frame.argtypes = C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    lib.ref_process_timed.restype, lib.ref_process_timed.argtypes = C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
    for f in frames:
        lib.ref_add_frame(ref.h, np.ascontiguousarray(f).ctypes.data)
    nm = C.c_int(0)
    sampler = ClockSampler(local)

    def step_e2e():
        for i in order:
            lib.ref_process_frame(ref.h, i, 1, 1, C.byref(nm))

    for _ in range(args.warmup):
        step_e2e()
    sampler.start()
    per = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        step_e2e()
        per.append((time.perf_counter() - t0) * 1e3)  # every call ends in cudaDeviceSynchronize: wall == device
    ms_e2e = sum(per)
    clocks = sampler.stop()
    # resident leg + breakdown: the same F distinct frames, every phase timed on its own
    ms4 = (C.c_double * 4)()
    acc = np.zeros(4)
    for it in range(args.warmup + args.steps):
        for i in order:
            lib.ref_process_timed(ref.h, i, 1, ms4)
            if it >= args.warmup:
                acc += np.array(list(ms4))
    total = F * args.steps
    h2d, detect, d2h, match = (acc / total).tolist()
    ms_res = (detect + match) * total
    per.sort()
    return dict(value=total / (ms_res * 1e-3), ms_per_step=ms_res / args.steps, e2e_value=total / (ms_e2e * 1e-3),
                e2e_ms_per_step_median=per[len(per) // 2], frames_per_step=F, clocks=clocks, kpts=int(lib.ref_total_size(ref.h)),
                breakdown={"h2d_upload": h2d, "detect": detect, "copy_to_host": d2h, "match": match},
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


def other_sizes(world):
    """BASELINE.json's other detector configurations: configs[1] 752x480, configs[3] 1600x900 (at 8 GPUs: the batch of
    256 frames, 32 per GPU, frame-parallel)."""
    out = [("752x480", 752, 480, 64), ("1600x900", 1600, 900, 32 if world == 8 else 64)]
    return [o for o in out if (o[1], o[2]) != (COLS, ROWS)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--frames", type=int, default=64, help="frames per step per GPU")
    ap.add_argument("--size", default="1920x1200", help="frame size WxH: 1920x1200 is BASELINE.json's metric; 752x480 and 1600x900 are its other configs")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline configuration only (no other sizes, no 100k x 100k matcher)")
    ap.add_argument("--quick", action="store_true", help="profiler runs: skip e2e leg, per-class profile, extras and CPU baselines")
    args = ap.parse_args()
    protect_stdout()
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
        cfg = workload_config(COLS, ROWS, args.frames, 1)   # exactly the keys and values of the other arm's `config`
        line = {
            "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "frames/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
            "run_detail": {"reference": "unmodified reference CUDA build (oracle/_ref) on 1 B200; the reference is single-GPU, so the "
                                        "same 1-GPU figure is reported for every N"},
            "e2e": {"value": r["e2e_value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "cpu_baseline": r["cpu_baseline"], "clocks": r.get("clocks"),
        }
        for k in ("breakdown", "kpts", "e2e_ms_per_step_median", "note"):
            if k in r:
                line["breakdown_ms_per_frame" if k == "breakdown" else k] = r[k]
        emit(line)
        return 0

    world, rank, local = dist_setup(args.gpus)
    m = measure_size(COLS, ROWS, args.frames, args.steps, args.warmup, world, rank, local, clocks=True, quick=args.quick)
    launches_total = sum_over_ranks(m["launches"], world)
    head = size_summary(m, world)
    clocks = m["clocks"]
    det = m["det"]
    # pure frame-graph time (no host sync per frame, no matcher) and per-class breakdown, rank 0 only
    ex = {}
    if rank == 0:
        import torch

        det.detectAndCompute(torch.from_numpy(m["frames"][0]).cuda())
        ex = dict(graph_ms=det.time_frames(5, 5 if args.quick else 100), pipe_ms=det.time_frames_pipelined(6, 6 if args.quick else 120),
                  prof=det.profile_frame(1 if args.quick else 3, 1 if args.quick else 20), kpts=det.total_size,
                  stages=det.time_stages(5 if args.quick else 50))
    fif = det.frames_in_flight
    main_frames = m["frames"]
    det.close()
    del m, det

    # the other BASELINE configurations and the sharded 100k x 100k matcher: every rank takes part (collective timing)
    extras, match100k = {}, None
    if not (args.quick or args.no_extras):
        for name, c, r_, f in other_sizes(world):
            mm = measure_size(c, r_, f, max(3, args.steps // 2), 3, world, rank, local)
            lat = None
            if rank == 0:
                import torch

                mm["det"].detectAndCompute(torch.from_numpy(mm["frames"][0]).cuda())
                lat = mm["det"].time_frames(5, 50)
            summ = size_summary(mm, world)
            summ["frame_graph_ms"] = lat
            if name == "1600x900" and world == 8:
                summ["note"] = "BASELINE.json configs[3]: batch of 256 frames, 32 per GPU, frame-parallel over 8 GPUs"
            extras[name] = summ
            mm["det"].close()
            del mm
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from bench_match import large_sharded

        match100k = large_sharded(100000, world, rank, iters=5, check_rows=64)

    if rank != 0:
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
            dist.destroy_process_group()
        return 0
    hbm_peak, _, peak_kind = measured_peaks()
    prof = ex["prof"]
    cb = class_bytes(ex["kpts"])
    # Roofline of the dominant kernel CLASS of the frame's critical path, the Gaussian-pyramid blur launches (blur_vh_kernel
    # for octave 0, blur_chain_kernel for the fused octaves: profiling classes blur_oct0 + blur_small). ONE method, the same in every round since r1's final commit: algorithmic
    # bytes of those launches (8 B per pixel per layer) over the CUDA-event time of the frame GRAPH
    # truncated after the pyramid (b200sift_time_stages: the launches exactly as the product runs them; the interval also
    # holds the single-CTA tail kernel, so the rate is a lower bound). The eager per-launch sum is printed beside it.
    groups = {"blur_vh_kernel+blur_chain_kernel": ["blur_oct0", "blur_small"], "detect_kernel": ["detect"], "descriptor_kernel": ["describe"]}
    gms = {g: sum(prof[k]["ms"] for k in ks) for g, ks in groups.items()}
    dom = "blur_vh_kernel+blur_chain_kernel"
    dom_ms = ex["stages"]["pyramid"]
    dom_bytes = sum(cb[k] for k in groups[dom])
    dom_launches = sum(prof[k]["launches"] for k in groups[dom])
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic = tj.get("blur", tj.get("blur_vh_kernel"))
    frame_bytes = 4 * COLS * ROWS + 48 * S_PIXELS + 284 * ex["kpts"]
    cfg = workload_config(COLS, ROWS, args.frames, world)   # the workload: identical keys and values in both arms
    detail = {"global_frames_per_step": args.frames * world, "parallelism": f"frames x{world}", "cpu_affinity": AFFINITY,
              "frames_in_flight_per_gpu": fif, "keypoints_per_frame": head["keypoints_per_frame"],
                "l2": f"device-resident inputs are {min(args.frames, DISTINCT_FRAMES)} distinct frames = {min(args.frames, DISTINCT_FRAMES) * COLS * ROWS * 4 / 1e6:.0f} MB; with the {fif} pyramids in flight "
                      f"({fif * 24 * S_PIXELS / 1e6:.0f} MB) the working set is {'larger' if (min(args.frames, DISTINCT_FRAMES) * COLS * ROWS * 4 + fif * 24 * S_PIXELS) > 126e6 else 'SMALLER'} than the 126 MB L2"}
    line = {
        "metric": METRIC, "value": head["value"], "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": head["ms_per_frame"] * args.frames, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg, "run_detail": detail,
        "ms_per_frame": head["ms_per_frame"],
        "ms_per_step_median": head["ms_per_step_median"], "ms_per_step_min": head["ms_per_step_min"], "ms_per_step_max": head["ms_per_step_max"],
        "frame_graph_ms": ex["graph_ms"],             # latency of one frame's graphs, nothing else in flight
        "frame_pipelined_ms": ex["pipe_ms"],          # per frame with every frame slot in flight (graphs only)
        "e2e": {"value": head["e2e"], "unit": "frames/s", "h2d_bytes_per_step": head["h2d_bytes_per_step"],
                "d2h_bytes_per_step": head["d2h_bytes_per_step"], "ms_per_frame": head["e2e_ms_per_frame"]},
        "e2e_u8": {"value": head["e2e_u8"], "unit": "frames/s", "h2d_bytes_per_step": args.frames * COLS * ROWS,
                   "note": "same public call with the 8-bit host frames the float images are cast from (b200sift_detect_host_u8_async): "
                           "identical results, a quarter of the H2D bytes; the float figure above is the headline"},
        "breakdown_ms_per_frame": head["breakdown_ms_per_frame"],
        "gpu_launches": int(launches_total),
        "clocks": None,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "peak_kind": peak_kind,
                     "launches_per_frame": dom_launches, "ms_per_frame": dom_ms,
                     "bytes_per_launch": dom_bytes / max(dom_launches, 1), "us_per_launch": dom_ms * 1e3 / max(dom_launches, 1),
                     "by_kernel_eager": {g: {"ms_per_frame": round(gms[g], 5),
                                             "GB/s": round(sum(cb[k] for k in groups[g]) / (gms[g] * 1e-3) / 1e9, 1)} for g in groups},
                     "eager_ms_per_frame": gms[dom],
                     "method": "algorithmic bytes of the pyramid launches / CUDA-event time of the frame graph truncated after the pyramid "
                               "(in-graph, as the product runs them; lower bound: the interval also holds the tail kernel); by_kernel_eager = "
                               "the same launches run eagerly on one stream with an event after each"},
        "roofline_frame": {"bytes": frame_bytes, "achieved": frame_bytes / (ex["pipe_ms"] * 1e-3) / 1e9,
                           "frac": frame_bytes / (ex["pipe_ms"] * 1e-3) / 1e9 / hbm_peak,
                           "note": "SURVEY 8(d) algorithmic bytes of one frame / pipelined frame time"},
        "kernel_ms_per_frame": {k: round(v["ms"], 5) for k, v in prof.items()},
        "graph_cumulative_ms": {k: round(v, 5) for k, v in ex["stages"].items()},
    }
    return line, args, world, main_frames, extras, match100k, clocks


def finish(line, args, world, main_frames, extras, match100k, clocks):
    line["clocks"] = clocks
    if extras:
        line["configs"] = extras
    if match100k is not None:
        line["match_100k"] = match100k
    if not args.quick:
        # second half of BASELINE.json's metric: latency of the 2000 x 2000 x 128 ratio-test match (seeded synthetic
        # descriptor sets, device-resident, median of per-call CUDA-event times) + its tensor roofline
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from bench_match import match_latency

        line["match_2k2k"] = match_latency(2000, iters=200)
    if world == 1 and not args.no_cpu_baseline and not args.quick:
        line["cpu_baseline"] = cpu_port_baseline(main_frames)
        line["opencv_cpu"] = opencv_baseline(main_frames)
        line["cpu_opencv_752x480"] = opencv_config0()
    emit(line)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    _r = main()
    sys.exit(_r if isinstance(_r, int) else finish(*_r))
