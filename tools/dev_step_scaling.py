"""Development measurement: how the time of one bench step depends on the number of frames per synchronisation, and how
much of it the host thread spends enqueueing (resident leg, Python loop over the C ABI)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from another_cuda_sift_b200 import CudaSiftConfig, Detector
from bench import make_frames, NUM_FEATURES

cols, rows = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "1920x1200").split("x"))
det = Detector(CudaSiftConfig(col_width=cols, row_width=rows, numFeatures=NUM_FEATURES))
assert det.gpuWarmUpAndAllocate()
base = make_frames(32, 0, rows, cols)
d_frames = [torch.from_numpy(f).cuda() for f in base]
idx = [torch.empty(NUM_FEATURES, dtype=torch.int32, device="cuda") for _ in range(det.frames_in_flight)]
for F in (16, 32, 64, 128, 256, 512):
    res = []
    for rep in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(F):
            det.detectAndComputeAsync(d_frames[i % 32])
            det.matchConsecutiveAsync(idx[i % len(idx)])
        t1 = time.perf_counter()
        det.synchronize()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        res.append(((t2 - t0) * 1e3, (t1 - t0) * 1e3))
    tot, host = min(r[0] for r in res[1:]), min(r[1] for r in res[1:])
    print(f"F={F:4d} step {tot:8.3f} ms  host enqueue {host:8.3f} ms  per frame {tot / F:.4f} ms  host per frame {host / F:.4f} ms  frames/s {F / tot * 1e3:.0f}", flush=True)
