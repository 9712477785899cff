"""Dev experiment (GPU box): throughput of one Detector vs two Detectors fed alternate frames (own streams)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from another_cuda_sift_b200 import CudaSiftConfig, Detector
from another_cuda_sift_b200.synth import synth_frame, next_frame

cols, rows = 1920, 1200
frames = []
img = synth_frame(rows, cols, 0)
for i in range(16):
    if i: img = next_frame(img, i)
    frames.append(torch.from_numpy(img).cuda())
dets = [Detector(CudaSiftConfig(col_width=cols, row_width=rows)) for _ in range(3)]
for d in dets: d.gpuWarmUpAndAllocate()

def run(n_det, iters=400):
    for k in range(20):
        dets[k % n_det].detectAndComputeAsync(frames[k % 16])
    for d in dets[:n_det]: d.synchronize()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(iters):
        dets[k % n_det].detectAndComputeAsync(frames[k % 16])
    for d in dets[:n_det]: d.synchronize()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"{n_det} detector(s): {dt / iters * 1e3:.4f} ms/frame, {iters / dt:.0f} frames/s")

for n in (1, 2, 3, 1, 2):
    run(n)
