"""Stage replay against a dump written by Detector.setDataGen / b200sift_dump_frame.

Counterpart of the reference's tool/perf.cu:17-113 (load one frame's serialized stage inputs + expected
outputs, run filter / resize / minus / findPeaks / adjust / orientation / descriptor, print one verdict per
stage), over raw arrays instead of msgpack+zlib. The engine fuses several of the reference's stages (DoG is
never stored, the gather is part of the descriptor kernel), so the stages replayed here are the engine's:

    pyramid      Gaussian stacks of every octave      (Filter + Resize of the reference)
    extrema      raw extrema count per octave         (Minus + findPeaks3D)
    refine       refined keypoints / features         (adjustExtrema)
    orientation  oriented keypoint lists              (calOriHistMultiThread + collectKpts)
    descriptor   final keypoints / features / descriptors (genDescriptor + collectAcrossOctaves)

    python tools/stage_replay.py DUMP_DIR                 # run every stage, compare, exit 1 on a mismatch
    python tools/stage_replay.py DUMP_DIR --stage refine  # stop after one stage (ncu: profile it in isolation)
    python tools/stage_replay.py --make DUMP_DIR [--cols 752 --rows 480 --seed 1234]   # write a dump of a synthetic frame

The engine is deterministic (fixed-point histograms), so a dump written by the same build replays bit-exactly;
across builds the descriptor stage is compared with the parity tolerance (normalised L2 <= 1e-3).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

STAGES = {"pyramid": 1, "extrema": 3, "refine": 4, "orientation": 5, "descriptor": 7}


def load_dump(path: str) -> dict:
    with open(os.path.join(path, "manifest.json")) as f:
        man = json.load(f)
    if man.get("format") != "b200sift-stage-dump-1":
        raise ValueError(f"{path}: not a b200sift stage dump")
    cfg = man["config"]

    def arr(name, dtype, shape):
        a = np.fromfile(os.path.join(path, name), dtype=dtype)
        return a.reshape(shape)

    d = {"manifest": man, "input": arr("input.f32", np.float32, (cfg["row_width"], cfg["col_width"])), "oct": []}
    for o, oc in enumerate(man["octaves"]):
        M, T = min(oc["refined"], oc["cap"]), min(oc["oriented"], oc["cap"])
        d["oct"].append({
            "gauss": arr(f"oct{o}_gauss.f32", np.float32, (6, oc["h"], oc["w"])),
            "refined_kpts": arr(f"oct{o}_refined_kpts.f32x3", np.float32, (M, 3)),
            "refined_features": arr(f"oct{o}_refined_features.f32x4", np.float32, (M, 4)),
            "oriented_kpts": arr(f"oct{o}_oriented_kpts.f32x3", np.float32, (T, 3)),
            "oriented_features": arr(f"oct{o}_oriented_features.f32x4", np.float32, (T, 4))})
    n = man["total_size"]
    if os.path.exists(os.path.join(path, "final_kpts.f32x3")):
        d["final_kpts"] = arr("final_kpts.f32x3", np.float32, (n, 3))
        d["final_features"] = arr("final_features.f32x4", np.float32, (n, 4))
        d["final_descriptors"] = arr("final_descriptors.f16x128", np.float16, (n, 128))
    return d


def same_bits(a: np.ndarray, b: np.ndarray) -> bool:
    return a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


def replay(path: str, last: str = "descriptor", verbose: bool = True) -> dict:
    """Returns {stage: bool}. Imports the product (CUDA) path only."""
    from another_cuda_sift_b200 import CudaSiftConfig, Detector

    d = load_dump(path)
    cfg = d["manifest"]["config"]
    det = Detector(CudaSiftConfig(col_width=cfg["col_width"], row_width=cfg["row_width"], numFeatures=cfg["numFeatures"],
                                  numOctaveLayers=cfg["numOctaveLayers"], contrastThreshould=cfg["contrastThreshould"],
                                  edgeThreshould=cfg["edgeThreshould"], sigma=cfg["sigma"]))
    assert det.gpuWarmUpAndAllocate()
    octs = d["manifest"]["octaves"]
    verdict = {}
    for name, code in STAGES.items():
        if code == 7:
            det.detectAndCompute(d["input"])
            det.copyToHost(True)
        else:
            det.debug_run_stages(d["input"], code)
        if name == "pyramid":
            ok = all(same_bits(det.debug_gauss(o), d["oct"][o]["gauss"]) for o in range(len(octs)))
        elif name == "extrema":
            ok = all(det.debug_counts(o)["P"] == octs[o]["raw_extrema"] for o in range(len(octs)))
        elif name == "refine":
            ok = True
            for o, oc in enumerate(octs):
                M = min(oc["refined"], oc["cap"])
                ok &= det.debug_counts(o)["M"] == oc["refined"]
                # the engine compacts the refined keypoints into coll[0, M) (the last slot is quirk Q5's stale one)
                ok &= same_bits(det.debug_kpt_buffer(o, "coll")[:M], d["oct"][o]["refined_kpts"])
                ok &= same_bits(det.debug_kpt_buffer(o, "collF")[:M], d["oct"][o]["refined_features"])
        elif name == "orientation":
            ok = True
            for o, oc in enumerate(octs):
                T = min(oc["oriented"], oc["cap"])
                ok &= det.debug_counts(o)["T"] == oc["oriented"]
                ok &= same_bits(det.debug_kpt_buffer(o, "cand")[:T], d["oct"][o]["oriented_kpts"])
                ok &= same_bits(det.debug_kpt_buffer(o, "candF")[:T], d["oct"][o]["oriented_features"])
        else:
            n = d["manifest"]["total_size"]
            ok = det.total_size == n and "final_kpts" in d
            if ok:
                ok &= same_bits(det.final_kpts[:n], d["final_kpts"]) and same_bits(det.final_features[:n], d["final_features"])
                a = det.descriptors[:n].astype(np.float32) / 512.0
                b = d["final_descriptors"].astype(np.float32) / 512.0
                err = np.sqrt(((a - b) ** 2).sum(axis=1)).max() if n else 0.0
                ok &= bool(err <= 1e-3)
                if verbose:
                    print(f"  descriptor max normalised L2 error {err:.2e} over {n} keypoints")
        verdict[name] = bool(ok)
        if verbose:
            print(f"{name}: {int(ok)}")
        if name == last:
            break
    det.close()
    return verdict


def make(path: str, cols: int, rows: int, seed: int, num_features: int) -> None:
    from another_cuda_sift_b200 import CudaSiftConfig, Detector
    from another_cuda_sift_b200.synth import synth_frame

    det = Detector(CudaSiftConfig(col_width=cols, row_width=rows, numFeatures=num_features))
    assert det.gpuWarmUpAndAllocate()
    det.setDataGen(path)
    det.detectAndCompute(synth_frame(rows, cols, seed - 1234))
    det.close()
    print(f"wrote {path}: {det.total_size} keypoints")


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("path", help="dump directory")
    ap.add_argument("--stage", default="descriptor", choices=list(STAGES))
    ap.add_argument("--make", action="store_true", help="write a dump of a synthetic frame instead of replaying one")
    ap.add_argument("--cols", type=int, default=752)
    ap.add_argument("--rows", type=int, default=480)
    ap.add_argument("--seed", type=int, default=1234, help="frame seed (1234 + frame index, SURVEY.md 8(d))")
    ap.add_argument("--num-features", type=int, default=5000)
    a = ap.parse_args()
    if a.make:
        make(a.path, a.cols, a.rows, a.seed, a.num_features)
        return 0
    v = replay(a.path, a.stage)
    return 0 if all(v.values()) else 1


if __name__ == "__main__":
    sys.exit(main())
