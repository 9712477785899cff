"""Profiling target: a few matchBruteForce calls (device API) at n x n. usage: prof_match.py [n] [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from another_cuda_sift_b200 import matchBruteForce
from another_cuda_sift_b200.synth import synth_descriptors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
q, t = synth_descriptors(n, n)
dq = torch.from_numpy(q.view(np.int16)).cuda(); dt = torch.from_numpy(t.view(np.int16)).cuda()
out = torch.empty(n, dtype=torch.int32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for _ in range(iters):
    matchBruteForce(dq, n, dt, n, stream=s, out=out)
torch.cuda.synchronize()
print("matches", int((out >= 0).sum()))
