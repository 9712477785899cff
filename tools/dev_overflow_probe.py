"""Development probe: a pure-noise frame drives octave 0 past its slot capacity (cap = 10 * numFeatures = 50,000 raw
extrema) at the default configuration; engine against the live reference build."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from another_cuda_sift_b200 import CudaSiftConfig, Detector
from oracle.pyoracle import RefOracle, ref_available
from util import check_results

cols, rows = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "1920x1200").split("x"))
assert ref_available()
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
det = Detector(CudaSiftConfig(col_width=cols, row_width=rows, numFeatures=nf)); assert det.gpuWarmUpAndAllocate()
ref = RefOracle(cols, rows, nf)
rng = np.random.default_rng(3)
for fr in range(3):
    g = rng.random((rows // 5 + 1, cols // 5 + 1)) < 0.5
    img = (np.kron(g, np.ones((5, 5)))[:rows, :cols] * 255).astype(np.float32)
    ref.detect(img)
    det.detectAndCompute(img)
    det.copyToHost(True)
    got = dict(n=det.total_size, kpts=det.final_kpts, feats=det.final_features, desc=det.descriptors)
    want = ref.results()
    print("frame", fr, "engine total", got["n"], "reference total", want["n"])
    print("  engine counts per octave (P, M, T):", [tuple(det.debug_counts(o).values()) for o in range(4)], "cap", det.octaves[0]["cap"])
    octs = [(o["w"], o["h"], o["pitch"], o["cap"]) for o in det.octaves]
    try:
        print("  ", check_results(got, want, octaves=octs, max_undefined=1.0))
    except AssertionError as e:
        print("  MISMATCH", str(e)[:600])
