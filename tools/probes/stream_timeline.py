"""Development probe (not product, not run by the tests): prints the clock64 stamps of one CTA of the streaming
matcher filter -- when the MMA-issuing thread starts waiting for a drained accumulator, passes the barrier and has
issued its 9 MMAs; when a scan warp starts waiting for the MMAs, sees them done and has released the accumulator.

It needs an INSTRUMENTED build of csrc/match_tc.cu: a `__device__ long long g_tl[1024]` written at those six points
(index step * 16 + {0,1,2} issuer half 0, {4,5,6} issuer half 1, {8,9,10} scan warp half 0, {12,13,14} half 1) for
one (blockIdx.x, blockIdx.y), and an exported `int b200sift_debug_timeline(long long* out)` that copies it out with
cudaMemcpyFromSymbol. The stock library does not export that symbol and this script then fails at the lookup.
profiles/r1_match_stream.md quotes what it showed."""
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
