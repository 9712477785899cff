"""Dev script (GPU box): stage-by-stage comparison engine vs oracle/_ref vs CPU oracle."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from another_cuda_sift_b200 import CudaSiftConfig, Detector, matchBruteForce
from another_cuda_sift_b200.synth import synth_frame, next_frame, synth_descriptors
from oracle.pyoracle import RefOracle, CpuOracle, ref_available

def cmp(name, a, b):
    a = np.asarray(a); b = np.asarray(b)
    if a.shape != b.shape:
        print(f"  {name}: SHAPE {a.shape} vs {b.shape}"); return False
    eq = np.array_equal(a.view(np.uint8) if a.dtype.kind=='f' else a, b.view(np.uint8) if b.dtype.kind=='f' else b)
    if eq:
        print(f"  {name}: bit-exact ({a.size} elems)")
    else:
        d = np.abs(a.astype(np.float64) - b.astype(np.float64))
        print(f"  {name}: DIFF max={d.max():.6g} n_diff={(d>0).sum()}/{a.size}")
    return eq

def run(cols, rows, nframes=3, cpu=True):
    print(f"=== {cols}x{rows}")
    det = Detector(CudaSiftConfig(col_width=cols, row_width=rows)); det.gpuWarmUpAndAllocate()
    ref = RefOracle(cols, rows, 5000)
    cpuo = CpuOracle(cols, rows, 5000) if cpu else None
    print(" octaves mine", det.octaves[:3], "ref", ref.octaves[:3])
    for i in range(-1, 6):
        cmp(f"taps[{i}]", det.debug_taps(i), ref.taps(i))
    img = synth_frame(rows, cols, 0)
    for fr in range(nframes):
        if fr > 0: img = next_frame(img, fr)
        print(f"-- frame {fr}")
        ref.set_image(img)
        # stage-wise on ref
        ref.stage(0); ref.stage(1)
        det.debug_run_stages(img, 1)
        for o in range(det.n_octaves):
            ok = cmp(f"gauss[{o}]", det.debug_gauss(o), ref.stack(o, "gauss"))
        ref.stage(2); ref.stage(3)
        det.debug_run_stages(img, 3)
        for o in range(det.n_octaves):
            P = ref.counts_raw(o)[0]
            c = det.debug_counts(o)
            n = min(P, ref.octaves[o]["cap"] - 1)
            print(f"  oct{o}: P ref={P} mine={c['P']}")
            cmp(f"cand[{o}][:{n}]", det.debug_kpt_buffer(o, "cand")[:n], ref.kpt_buffer(o, "cand")[:n])
        ref.stage(4)
        det.debug_run_stages(img, 4)
        for o in range(det.n_octaves):
            M = ref.counts_raw(o)[1]; c = det.debug_counts(o)
            print(f"  oct{o}: M ref={M} mine={c['M']}")
            n = max(M - 1, 0)
            cmp(f"coll[{o}][:{n}]", det.debug_kpt_buffer(o, "coll")[:n], ref.kpt_buffer(o, "coll")[:n])
            cmp(f"collF[{o}][:{n}].xyz", det.debug_kpt_buffer(o, "collF")[:n, :3], ref.kpt_buffer(o, "collF")[:n, :3])
        ref.stage(5)
        det.debug_run_stages(img, 5)
        for o in range(det.n_octaves):
            T = ref.counts_raw(o)[1]; c = det.debug_counts(o)
            print(f"  oct{o}: T ref={T} mine={c['T']}")
            n = max(min(T, c['T']) - 1, 0)
            cmp(f"candO[{o}][:{n}]", det.debug_kpt_buffer(o, "cand")[:n], ref.kpt_buffer(o, "cand")[:n])
            cmp(f"candFO[{o}][:{n}]", det.debug_kpt_buffer(o, "candF")[:n], ref.kpt_buffer(o, "candF")[:n])
        ref.stage(6); ref.stage(7)
        det.debug_run_stages(img, 7)
        det.copyToHost(True)
        r = ref.results()
        n = min(r["n"], det.total_size)
        print(f"  total ref={r['n']} mine={det.total_size}")
        cmp("final kpts", det.final_kpts[:n], r["kpts"][:n])
        cmp("final feats", det.final_features[:n], r["feats"][:n])
        a = det.descriptors[:n].astype(np.float32); b = r["desc"][:n].astype(np.float32)
        dd = np.abs(a - b)
        l2 = np.linalg.norm((a - b) / 512.0, axis=1)
        print(f"  desc: max|d|={dd.max() if n else 0:.3f} rows>1: {(dd.max(axis=1)>1).sum() if n else 0}/{n}  L2 max={l2.max() if n else 0:.5f} mean={l2.mean() if n else 0:.6f}")
        if cpuo is not None:
            nc = cpuo.detect(img); rc = cpuo.results()
            print(f"  cpu oracle total={nc}")
            m = min(nc, r["n"])
            cmp("cpu-vs-ref kpts", rc["kpts"][:m], r["kpts"][:m])
            cmp("cpu-vs-ref feats", rc["feats"][:m], r["feats"][:m])
            for o in range(min(3, det.n_octaves)):
                cmp(f"cpu-vs-ref gauss[{o}]", cpuo.stack(o, "gauss"), ref.stack(o, "gauss"))
                print("   cpu counts", cpuo.counts(o))
    # graph path twice + timing
    det2 = Detector(CudaSiftConfig(col_width=cols, row_width=rows)); det2.gpuWarmUpAndAllocate()
    img0 = synth_frame(rows, cols, 0)
    det2.detectAndCompute(img0); det2.copyToHost(True)
    print(" graph path total", det2.total_size, "launches", det2.launches_per_frame())
    print(" mine ms/frame (device resident):", det2.time_frames(5, 50))
    ref2 = RefOracle(cols, rows, 5000); ref2.set_image(img0)
    for _ in range(3): ref2._detect(ref2.h)
    print(" ref ms/frame device-resident:", ref2.time_device_resident(3, 20), " e2e:", ref2.time_e2e(3, 20, True))

if __name__ == "__main__":
    run(752, 480)
    run(1920, 1200, nframes=2, cpu=False)
    q, t = synth_descriptors(2000, 2000)
    mine = matchBruteForce(q, 2000, t, 2000)
    ref = RefOracle(64, 64, 100)
    r = ref.match(q, t)
    print("match 2k: equal", np.array_equal(mine, r), "matches", (r >= 0).sum(), "diff", (mine != r).sum())
    cpu = CpuOracle(64, 64, 100)
    c = cpu.match(q[:300], t)
    print("cpu-vs-ref match[:300] equal", np.array_equal(c, r[:300]))
    print("ref match ms:", ref.time_match(q, t, 3, 20))
    import torch
    dq = torch.from_numpy(q.view(np.int16)).cuda(); dt = torch.from_numpy(t.view(np.int16)).cuda()
    out = torch.empty(2000, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(3): matchBruteForce(dq, 2000, dt, 2000, stream=s, out=out)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): matchBruteForce(dq, 2000, dt, 2000, stream=s, out=out)
    e1.record(); torch.cuda.synchronize()
    print("mine match ms:", e0.elapsed_time(e1) / 20, "equal", np.array_equal(out.cpu().numpy(), r))
