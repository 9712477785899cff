"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libref_sift.so, built from
/root/reference by oracle/Makefile) running on a B200.  Run on the GPU box:

    python tests/golden/make_golden.py [outdir]

The reference ships no golden vectors: these files are what pins the CPU oracle
(tests/test_oracle_golden.py, CPU-only) and the CUDA engine (tests/test_gpu_golden.py).  Inputs are
regenerated from seeds by another_cuda_sift_b200.synth, and stored as uint8 too so a generator
drift is detected rather than silently followed.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from another_cuda_sift_b200.synth import blocks_frame, next_frame, synth_descriptors, synth_frame  # noqa: E402
from oracle.pyoracle import RefOracle  # noqa: E402


def run_frames(cols, rows, nframes, keep_stacks=0, density=1.0, num_features=5000):
    """Stage-by-stage capture of `nframes` consecutive frames on ONE reference Detector (frame history
    matters for the stale slots of quirk Q5)."""
    ref = RefOracle(cols, rows, num_features)
    out = {"cols": cols, "rows": rows, "nframes": nframes, "n_octaves": ref.n_octaves, "density": density,
           "num_features": num_features,
           "octaves": np.array([[o["w"], o["h"], o["pitch"], o["cap"]] for o in ref.octaves], np.int32)}
    img = synth_frame(rows, cols, 0, density=density)
    for fr in range(nframes):
        if fr > 0:
            img = next_frame(img, fr)
        if cols * rows <= 64 * 1024:
            out[f"f{fr}_img_u8"] = img.astype(np.uint8)
        else:  # big frames: only a hash; tests regenerate them from the seed and verify it
            out[f"f{fr}_img_sha1"] = np.array(hashlib.sha1(img.astype(np.float32).tobytes()).hexdigest())
        ref.set_image(img)
        ref.stage(0)
        ref.stage(1)
        for o in range(min(keep_stacks, ref.n_octaves) if fr == 0 else 0):  # full stacks for frame 0 only
            out[f"f{fr}_gauss{o}"] = ref.stack(o, "gauss")
        # cheap whole-pyramid pin: float64 sum and xor of the bit patterns of every Gaussian layer
        sums, xors = [], []
        for o in range(ref.n_octaves):
            g = ref.stack(o, "gauss")
            sums.append(g.astype(np.float64).sum(axis=(1, 2)))
            xors.append(np.bitwise_xor.reduce(g.view(np.uint32).reshape(6, -1), axis=1))
        out[f"f{fr}_gauss_sum"] = np.array(sums)
        out[f"f{fr}_gauss_xor"] = np.array(xors, np.uint32)
        ref.stage(2)
        ref.stage(3)
        P = []
        for o in range(ref.n_octaves):
            p = ref.counts_raw(o)[0]
            P.append(p)
            n = min(p, ref.octaves[o]["cap"] - 1)
            out[f"f{fr}_extrema{o}"] = ref.kpt_buffer(o, "cand")[:n].astype(np.int16)
        ref.stage(4)
        M = []
        for o in range(ref.n_octaves):
            m = ref.counts_raw(o)[1]
            M.append(m)
            out[f"f{fr}_refined{o}"] = ref.kpt_buffer(o, "coll")[:max(m - 1, 0)]
            out[f"f{fr}_refinedF{o}"] = ref.kpt_buffer(o, "collF")[:max(m - 1, 0), :3]
        ref.stage(5)
        T = []
        for o in range(ref.n_octaves):
            t = ref.counts_raw(o)[1]
            T.append(t)
            out[f"f{fr}_oriented{o}"] = ref.kpt_buffer(o, "cand")[:max(t - 1, 0)]
            out[f"f{fr}_orientedF{o}"] = ref.kpt_buffer(o, "candF")[:max(t - 1, 0)]
        ref.stage(6)
        ref.stage(7)
        r = ref.results()
        n = r["n"]
        out[f"f{fr}_counts"] = np.array([P, M, T], np.int32)
        out[f"f{fr}_total"] = n
        out[f"f{fr}_kpts"] = r["kpts"][:n]
        out[f"f{fr}_feats"] = r["feats"][:n]
        out[f"f{fr}_desc"] = r["desc"][:n]
    ref.close()
    return out


def envelope(cols, rows, runs=3, kind="texture"):
    """The reference's own run-to-run spread (float shared-memory atomics in orientation / descriptor)."""
    img = blocks_frame(rows, cols, 5) if kind == "blocks" else synth_frame(rows, cols, 0)
    res = []
    for _ in range(runs):
        ref = RefOracle(cols, rows, 5000)
        ref.detect(img)
        res.append(ref.results())
        ref.close()
    n = [r["n"] for r in res]
    m = min(n)
    dmax, l2max = 0.0, 0.0
    same_kp = True
    for r in res[1:]:
        same_kp &= bool(np.array_equal(r["kpts"][:m], res[0]["kpts"][:m]) and np.array_equal(r["feats"][:m], res[0]["feats"][:m]))
        d = r["desc"][:m].astype(np.float32) - res[0]["desc"][:m].astype(np.float32)
        dmax = max(dmax, float(np.abs(d).max()))
        l2max = max(l2max, float(np.linalg.norm(d / 512.0, axis=1).max()))
    return dict(totals=np.array(n), same_keypoints=same_kp, desc_max_abs=dmax, desc_l2_max=l2max)


def envelopes(outdir):
    """Run-to-run envelopes of the reference on the frames where the descriptor bar is asserted: the 752x480 texture,
    the axis-aligned blocks frame (tests/test_gpu_parity.py) and 1920x1200."""
    for name, args in (("envelope_752x480", (752, 480)), ("envelope_blocks_640x360", (640, 360, 3, "blocks")),
                       ("envelope_1920x1200", (1920, 1200))):
        env = envelope(*args)
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **env)
        print(name, {k: (v.tolist() if hasattr(v, "tolist") else v) for k, v in env.items()})


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    outdir = args[0] if args else os.path.dirname(os.path.abspath(__file__))
    os.makedirs(outdir, exist_ok=True)
    if "--envelopes-only" in sys.argv:
        return envelopes(outdir)
    ref = RefOracle(64, 64, 100)
    taps = {f"taps{i}": ref.taps(i) for i in range(-1, 6)}
    np.savez_compressed(os.path.join(outdir, "taps.npz"), **taps)

    np.savez_compressed(os.path.join(outdir, "frames_160x120.npz"), **run_frames(160, 120, 3, keep_stacks=9))
    np.savez_compressed(os.path.join(outdir, "frames_203x151.npz"), **run_frames(203, 151, 2, keep_stacks=9))  # odd dims: Q3
    np.savez_compressed(os.path.join(outdir, "frames_752x480.npz"), **run_frames(752, 480, 3))
    np.savez_compressed(os.path.join(outdir, "frames_1920x1200.npz"), **run_frames(1920, 1200, 2))
    # tiny numFeatures (cap = 2000 slots): exercises the orientation slot cap (Q6) and the numFeatures truncation (Q7)
    np.savez_compressed(os.path.join(outdir, "frames_caps_752x480.npz"), **run_frames(752, 480, 2, num_features=200))

    q, t = synth_descriptors(2000, 2000)
    m = {"m2000": ref.match(q, t)}
    for nq, nt in ((5, 1), (7, 2), (64, 33), (100, 31), (33, 64), (257, 1000)):
        m[f"m{nq}x{nt}"] = ref.match(q[:nq], t[:nt])
    m["self300"] = ref.match(q[:300], q[:300])
    np.savez_compressed(os.path.join(outdir, "match.npz"), **m)

    envelopes(outdir)
    print("golden written to", outdir)


if __name__ == "__main__":
    main()
