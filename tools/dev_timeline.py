"""Development aid: steady-state schedule of the frame graphs with every slot in flight (b200sift_debug_timeline):
when each frame's stream reached part A and finished parts A (pyramid + extrema), B (ordered candidate stages) and
C (descriptors). Usage: python tools/dev_timeline.py [WxH] [frames] [slots]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from another_cuda_sift_b200 import CudaSiftConfig, Detector, _lib  # noqa: E402
from another_cuda_sift_b200.synth import synth_frame  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "1920x1200"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
slots = int(sys.argv[3]) if len(sys.argv) > 3 else 0
cols, rows = (int(v) for v in size.split("x"))
det = Detector(CudaSiftConfig(col_width=cols, row_width=rows), frames_in_flight=slots)
assert det.gpuWarmUpAndAllocate()
det.detectAndCompute(synth_frame(rows, cols, 0))
out = (C.c_float * (4 * n))()
_lib.check(det.lib.b200sift_debug_timeline(det._h, n, out))
t = np.array(list(out)).reshape(n, 4) * 1e3  # us
print(f"{size}, {det.frames_in_flight} frames in flight, {det.total_size} keypoints; times in us since the common start")
print("frame  reach_A   end_A   end_B   end_C |  A dur  wait+B  C dur | end_B - prev end_B   end_C - prev end_C")
for i in range(n):
    pb = t[i, 2] - t[i - 1, 2] if i else 0.0
    pc = t[i, 3] - t[i - 1, 3] if i else 0.0
    print(f"{i:5d} {t[i,0]:8.1f} {t[i,1]:7.1f} {t[i,2]:7.1f} {t[i,3]:7.1f} | {t[i,1]-t[i,0]:6.1f} {t[i,2]-t[i,1]:7.1f} {t[i,3]-t[i,2]:6.1f} | {pb:8.1f} {pc:18.1f}")
lo, hi = n // 4, 3 * n // 4   # the middle half: no ramp-up, no drain
bstart = np.maximum(t[1:, 1], t[:-1, 2])
bdur = t[1:, 2] - bstart
print(f"middle half: {np.diff(t[lo:hi, 3]).mean():.1f} us per frame; A {np.mean(t[lo:hi,1]-t[lo:hi,0]):.1f} us, "
      f"B (from max(own A end, previous B end)) {bdur[lo:hi].mean():.1f} us, C {np.mean(t[lo:hi,3]-t[lo:hi,2]):.1f} us, "
      f"B waited for its own A in {(t[1:,1] > t[:-1,2])[lo:hi].mean()*100:.0f}% of the frames, "
      f"frames between reach_A and end_C: {np.mean(t[lo:hi,3]-t[lo:hi,0]) / np.diff(t[lo:hi, 3]).mean():.1f}")
