"""Dev script (GPU box): per-octave P/M/T counts of one synthetic frame. usage: python tools/octave_counts.py [cols rows]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from another_cuda_sift_b200 import CudaSiftConfig, Detector
from another_cuda_sift_b200.synth import synth_frame

cols, rows = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1200)
det = Detector(CudaSiftConfig(col_width=cols, row_width=rows))
det.gpuWarmUpAndAllocate()
det.detectAndCompute(synth_frame(rows, cols, 0))
det.copyToHost(True)
print("total", det.total_size)
for o in range(det.n_octaves):
    print(o, det.octaves[o], det.debug_counts(o))
