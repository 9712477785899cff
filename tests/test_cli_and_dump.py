"""Command-line tools (reference tool/detection_example.cc, tool/extract_and_match_example.cc) and the
stage dump / replay (8(f)3; reference Detector::setDataGen + tool/perf.cu)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

LIBDIR = os.path.join(ROOT, "another_cuda_sift_b200")
sys.path.insert(0, os.path.join(ROOT, "tools"))


def build_tool(name: str, out_dir) -> str:
    exe = os.path.join(str(out_dir), name)
    subprocess.check_call(["g++", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tools", "cli"),
                           "-I", "/usr/local/cuda/include", os.path.join(ROOT, "tools", "cli", name + ".cc"), "-L", LIBDIR, "-lb200sift",
                           "-L", "/usr/local/cuda/lib64", "-lcudart", "-Wl,-rpath," + LIBDIR, "-o", exe])
    return exe


def write_pgm(path, img):
    with open(path, "wb") as f:
        f.write(b"P5\n# synthetic\n%d %d\n255\n" % (img.shape[1], img.shape[0]))
        f.write(np.clip(img, 0, 255).astype(np.uint8).tobytes())


def test_tools_compile_and_report_missing_inputs_like_the_reference(tmp_path):
    det = build_tool("detection_example", tmp_path)
    em = build_tool("extract_and_match_example", tmp_path)
    r = subprocess.run([det, "-p", str(tmp_path / "nope.pgm")], capture_output=True, text=True)
    assert r.returncode == 1 and "Image DNE" in r.stdout              # detection_example.cc:27-30
    r = subprocess.run([em, "--dir", str(tmp_path / "nope")], capture_output=True, text=True)
    assert r.returncode == 1 and "Directory does not exists." in r.stdout  # extract_and_match_example.cc:46-49


@pytest.mark.gpu
def test_tools_run_on_pgm_frames_and_agree_with_the_python_mirror(tmp_path):
    from another_cuda_sift_b200 import CudaSiftConfig, Detector, matchBruteForce
    from another_cuda_sift_b200.synth import next_frame, synth_frame

    frames = [synth_frame(240, 320, 0)]
    frames.append(next_frame(frames[0], 1))
    d = tmp_path / "imgs"
    d.mkdir()
    for i, f in enumerate(frames):
        write_pgm(d / f"f{i}.pgm", f)
    out = tmp_path / "out"
    out.mkdir()
    dump = tmp_path / "dump"
    dump.mkdir()
    r = subprocess.run([build_tool("detection_example", tmp_path), "-p", str(d / "f0.pgm"), "-d", str(dump), "-n", "3"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "r: 240, c: 320" in r.stdout and os.path.exists(dump / "manifest.json")
    r = subprocess.run([build_tool("extract_and_match_example", tmp_path), "-d", str(d), "-o", str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr

    det = Detector(CudaSiftConfig(col_width=320, row_width=240, numFeatures=2000))
    assert det.gpuWarmUpAndAllocate()
    det.detectAndCompute(frames[0]); det.copyToHost(True)
    n0, k0, d0 = det.total_size, det.final_kpts.copy(), det.descriptors.copy()
    det.detectAndCompute(frames[1]); det.copyToHost(True)
    n1 = det.total_size
    want = matchBruteForce(d0, n0, det.descriptors, n1)
    got = np.loadtxt(out / "matches_1.txt").reshape(-1, 6)
    keep = np.nonzero(want >= 0)[0]
    assert len(keep) > 10
    assert np.array_equal(got[:, 0].astype(int), keep) and np.array_equal(got[:, 1].astype(int), want[keep])
    assert np.allclose(got[:, 2:4], k0[keep, :2]) and np.allclose(got[:, 4:6], det.final_kpts[want[keep], :2])
    assert os.path.getsize(out / "matches_1.ppm") > 640 * 240 * 3
    assert json.load(open(dump / "manifest.json"))["total_size"] == n0


@pytest.mark.gpu
def test_stage_dump_replays_bit_exactly(tmp_path):
    import stage_replay
    from another_cuda_sift_b200 import CudaSiftConfig, Detector
    from another_cuda_sift_b200.synth import synth_frame

    path = str(tmp_path / "dump")
    det = Detector(CudaSiftConfig(col_width=376, row_width=240, numFeatures=3000))
    assert det.gpuWarmUpAndAllocate()
    det.setDataGen(path)
    img = synth_frame(240, 376, 5)
    det.detectAndCompute(img)
    det.copyToHost(True)
    d = stage_replay.load_dump(path)
    man = d["manifest"]
    assert man["total_size"] == det.total_size > 50 and len(man["octaves"]) == det.n_octaves
    assert np.array_equal(d["input"], img)
    assert np.array_equal(d["final_kpts"], det.final_kpts[: det.total_size])
    assert np.array_equal(d["final_descriptors"].view(np.uint16), det.descriptors[: det.total_size].view(np.uint16))
    assert sum(min(o["oriented"], o["cap"]) for o in man["octaves"]) >= det.total_size
    det.close()
    verdict = stage_replay.replay(path, verbose=False)
    assert verdict == {"pyramid": True, "extrema": True, "refine": True, "orientation": True, "descriptor": True}
    # a corrupted expected file must be noticed
    g = np.fromfile(os.path.join(path, "oct1_gauss.f32"), np.float32)
    g[7] += 1.0
    g.tofile(os.path.join(path, "oct1_gauss.f32"))
    assert stage_replay.replay(path, "pyramid", verbose=False) == {"pyramid": False}
