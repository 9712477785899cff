import sys, ctypes as C, numpy as np, torch
sys.path.insert(0, '/root/repo')
from another_cuda_sift_b200 import matchBruteForce, _lib
from another_cuda_sift_b200.synth import synth_descriptors
n = 100000
q, t = synth_descriptors(n, n)
dq = torch.from_numpy(q.view(np.int16)).cuda(); dt = torch.from_numpy(t.view(np.int16)).cuda()
out = torch.empty(n, dtype=torch.int32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for _ in range(2): matchBruteForce(dq, n, dt, n, stream=s, out=out)
torch.cuda.synchronize()
lib = _lib.load()
buf = (C.c_longlong * 1024)()
C.CDLL(lib._name).b200sift_debug_timeline(buf)
a = np.array(buf[:], dtype=np.int64).reshape(64, 16)
t0 = a[0, 0]
print("step | issuer h0: wait_begin drained issued | h1: ... | scanner h0: wait_begin mma_done arrived | h1")
for sidx in range(0, 50):
    r = a[sidx] - t0
    print(sidx, r[0:3], r[4:7], r[8:11], r[12:15])
