"""Dev script (GPU box): tensor-core matcher vs CPU oracle / CUDA-core path, + timings."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from another_cuda_sift_b200 import matchBruteForce, _lib
from another_cuda_sift_b200.synth import synth_descriptors
from oracle.pyoracle import CpuOracle

lib = _lib.load()
cpu = CpuOracle(64, 64, 100)

def dist_check(nq, nt):
    q, t = synth_descriptors(max(nq, 8), max(nt, 8)); q, t = q[:nq], t[:nt]
    dq = torch.from_numpy(q.view(np.int16)).cuda(); dt = torch.from_numpy(t.view(np.int16)).cuda()
    dd = torch.zeros(nq, nt, dtype=torch.float32, device="cuda"); di = torch.empty(nq, dtype=torch.int32, device="cuda")
    rc = lib.b200sift_debug_match_dist(dq.data_ptr(), nq, dt.data_ptr(), nt, dd.data_ptr(), di.data_ptr())
    assert rc == 0, lib.b200sift_last_error()
    a = q.astype(np.float64); b = t.astype(np.float64)
    D = (a * a).sum(1)[:, None] + (b * b).sum(1)[None, :] - 2 * a @ b.T
    got = dd.cpu().numpy()
    err = np.abs(got - D) / np.maximum(D, 1.0)
    want = cpu.match(q, t)
    print(f"dist {nq}x{nt}: max rel err {err.max():.3e}  idx equal {np.array_equal(di.cpu().numpy(), want)} ndiff {(di.cpu().numpy() != want).sum()}")
    return err.max()

def timeit(nq, nt, path, iters=20):
    lib.b200sift_match_set_path(path)
    q, t = synth_descriptors(nq, nt)
    dq = torch.from_numpy(q.view(np.int16)).cuda(); dt = torch.from_numpy(t.view(np.int16)).cuda()
    out = torch.empty(nq, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(3): matchBruteForce(dq, nq, dt, nt, stream=s, out=out)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): matchBruteForce(dq, nq, dt, nt, stream=s, out=out)
    e1.record(); torch.cuda.synchronize()
    lib.b200sift_match_set_path(0)
    return e0.elapsed_time(e1) / iters, out.cpu().numpy()

if __name__ == "__main__":
    for nq, nt in ((5, 3), (128, 128), (100, 257), (300, 500), (257, 4100)):
        dist_check(nq, nt)
    for n in (2000, 4400, 20000):
        t_tc, o_tc = timeit(n, n, 0, 20 if n < 10000 else 3)
        t_sm, o_sm = timeit(n, n, 1, 20 if n < 10000 else 3)
        print(f"{n}x{n}: tensor-core {t_tc:.4f} ms, cuda-core {t_sm:.4f} ms, equal {np.array_equal(o_tc, o_sm)} ndiff {(o_tc != o_sm).sum()} matches {(o_sm >= 0).sum()}")
