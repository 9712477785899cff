"""Dev probe: which duplicate keypoints does remove_duplicates miss (vs a host-side dedup of the default output)?"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from another_cuda_sift_b200 import CudaSiftConfig, Detector  # noqa: E402
from another_cuda_sift_b200.synth import next_frame, synth_frame  # noqa: E402

cols, rows = 752, 480
base = Detector(CudaSiftConfig(col_width=cols, row_width=rows)); base.gpuWarmUpAndAllocate()
dd = Detector(CudaSiftConfig(col_width=cols, row_width=rows), remove_duplicates=True); dd.gpuWarmUpAndAllocate()
img = synth_frame(rows, cols, 0)
for fr in range(3):
    if fr:
        img = next_frame(img, fr)
    out = []
    for d in (base, dd):
        d.detectAndCompute(img); d.copyToHost(True); n = d.total_size
        out.append((n, d.final_kpts[:n].copy(), d.final_features[:n].copy()))
        print(fr, "T per octave", [d.debug_counts(o)["T"] for o in range(d.n_octaves)])
    (na, ka, fa), (nb, kb, fb) = out
    key = lambda k, f: [(float(k[i, 0]), float(k[i, 1]), float(f[i, 1]), float(f[i, 3]), int(f[i, 0]) & 255) for i in range(len(k))]
    A, B = key(ka, fa), key(kb, fb)
    seen, dups = {}, []
    for i, t in enumerate(A):
        if t in seen:
            dups.append((i, seen[t], t, float(ka[i, 2]), float(fa[i, 0]), float(ka[seen[t], 2]), float(fa[seen[t], 0])))
        else:
            seen[t] = i
    print(fr, "default", na, "dedup", nb, "host-side dups", len(dups))
    from collections import Counter
    cb = Counter(B)
    for d in dups:
        print("   dup idx", d[0], "first idx", d[1], "key", d[2], "layer/packed", d[3:], "copies left in engine output:", cb[d[2]])
