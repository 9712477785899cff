"""Profiling driver: a few warm-up frames, then ONE 1920x1200 frame (+ consecutive-frame match) between
cudaProfilerStart / cudaProfilerStop, so that `ncu --profile-from-start off` sees exactly the launches of one frame
as the product runs them (frame graphs + matcher).

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file x.csv python tools/prof_frame.py
    ncu --profile-from-start off --set full --clock-control none --import-source on -o x python tools/prof_frame.py

--match N: instead, profile one N x N device matcher call (the streaming filter for N > 8192).
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from another_cuda_sift_b200 import CudaSiftConfig, Detector, _lib, matchBruteForce  # noqa: E402
from another_cuda_sift_b200.synth import next_frame, synth_descriptors, synth_frame  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--size", default="1920x1200")
ap.add_argument("--match", type=int, default=0)
ap.add_argument("--path", type=int, default=0)
args = ap.parse_args()
lib = _lib.load()
lib.b200sift_match_set_path(args.path)
if args.match:
    n = args.match
    q, t = synth_descriptors(n, n)
    dq, dt = torch.from_numpy(q.view(np.int16)).cuda(), torch.from_numpy(t.view(np.int16)).cuda()
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        out = matchBruteForce(dq, n, dt, n, stream=s)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    out = matchBruteForce(dq, n, dt, n, stream=s)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("matches", int((out >= 0).sum()))
    sys.exit(0)
cols, rows = (int(v) for v in args.size.split("x"))
det = Detector(CudaSiftConfig(col_width=cols, row_width=rows))
assert det.gpuWarmUpAndAllocate()
frames = [synth_frame(rows, cols, 0)]
for i in range(1, 4):
    frames.append(next_frame(frames[-1], i))
dev = [torch.from_numpy(f).cuda() for f in frames]
idx = torch.empty(5000, dtype=torch.int32, device="cuda")
for i in range(3):
    det.detectAndComputeAsync(dev[i])
    det.matchConsecutiveAsync(idx)
    det.synchronize()
torch.cuda.synchronize()
torch.cuda.profiler.start()
det.detectAndComputeAsync(dev[3])
det.matchConsecutiveAsync(idx)
det.synchronize()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("keypoints", det.total_size, "launches/frame", det.launches_per_frame())
