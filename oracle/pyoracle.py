"""TEST INFRASTRUCTURE -- ctypes front-end for the two checkers. NOT the product.

* ``CpuOracle``  -> oracle/liboracle_sift.so  (sift_oracle.cpp, CPU restatement)
* ``RefOracle``  -> oracle/_ref/libref_sift.so (the unmodified reference CUDA build; needs a GPU)

Both expose the same stage-by-stage view of `Detector::detectAndCompute`
(/root/reference/sift_cuda/interface/Detector.cu:133-233) so tests can compare the B200
engine against either of them buffer by buffer. Only tests/, __graft_entry__.smoke() and
bench.py's baseline legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CPU_LIB = os.path.join(HERE, "liboracle_sift.so")
REF_LIB = os.path.join(HERE, "_ref", "libref_sift.so")

STAGES = ("upload", "gaussian", "dog", "extrema", "refine", "orientation", "descriptor", "gather")


def build_cpu(force: bool = False) -> str:
    src = os.path.join(HERE, "sift_oracle.cpp")
    if force or not os.path.exists(CPU_LIB) or os.path.getmtime(CPU_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "cpu"], stdout=subprocess.DEVNULL)
    return CPU_LIB


def build_ref() -> str | None:
    """Build oracle/_ref from /root/reference when it is present (dev container only)."""
    if os.path.isdir("/root/reference/sift_cuda"):
        subprocess.check_call(["make", "-C", HERE, "ref", "-j8"], stdout=subprocess.DEVNULL)
    return REF_LIB if os.path.exists(REF_LIB) else None


def ref_available() -> bool:
    return os.path.exists(REF_LIB)


class _Oracle:
    prefix = ""
    lib_path = ""

    def __init__(self, cols: int, rows: int, num_features: int = 5000, num_octave_layers: int = 3,
                 contrast: float = 0.04, edge: float = 10.0, sigma: float = 1.6):
        self.lib = C.CDLL(self.lib_path)
        self.cols, self.rows, self.num_features = cols, rows, num_features
        self.n_layers = num_octave_layers + 3
        f = self._fn
        f("create", C.c_void_p, [C.c_int, C.c_int, C.c_int])
        f("create_cfg", C.c_void_p, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double])
        f("destroy", None, [C.c_void_p])
        f("num_octaves", C.c_int, [C.c_void_p])
        f("octave_info", None, [C.c_void_p, C.c_int, C.c_void_p])
        f("set_image", None, [C.c_void_p, C.c_void_p])
        f("stage", C.c_int, [C.c_void_p, C.c_int])
        f("detect", C.c_int, [C.c_void_p])
        f("total_size", C.c_int, [C.c_void_p])
        f("results", C.c_int, [C.c_void_p] * 5)
        f("get_stack", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p])
        f("get_kpt_buffer", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int])
        f("get_counts", None, [C.c_void_p, C.c_int, C.c_void_p])
        f("get_taps", C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int])
        f("match", C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p])
        if getattr(self, "_opts", 0):
            fn = self.lib.oracle_create_opts
            fn.restype = C.c_void_p
            fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
            self.h = fn(cols, rows, num_features, num_octave_layers, contrast, edge, sigma, self._opts)
        elif (num_octave_layers, contrast, edge, sigma) == (3, 0.04, 10.0, 1.6):
            self.h = self._create(cols, rows, num_features)
        else:
            self.h = self._create_cfg(cols, rows, num_features, num_octave_layers, contrast, edge, sigma)
        if not self.h:
            raise RuntimeError("oracle create failed")
        self.n_octaves = self._num_octaves(self.h)
        self.octaves = []
        for o in range(self.n_octaves):
            d = (C.c_int * 4)()
            self._octave_info(self.h, o, d)
            self.octaves.append(dict(w=d[0], h=d[1], pitch=d[2], cap=d[3]))

    def _fn(self, name, restype, argtypes):
        fn = getattr(self.lib, self.prefix + name)
        fn.restype, fn.argtypes = restype, argtypes
        setattr(self, "_" + name, fn)

    def close(self):
        if getattr(self, "h", None):
            self._destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- frame level
    def set_image(self, img: np.ndarray):
        img = np.ascontiguousarray(img, dtype=np.float32)
        assert img.shape == (self.rows, self.cols)
        self._set_image(self.h, img.ctypes.data)

    def stage(self, s: int):
        rc = self._stage(self.h, s)
        if rc != 0:
            raise RuntimeError(f"stage {s} failed: {rc}")

    def detect(self, img: np.ndarray) -> int:
        self.set_image(img)
        return self._detect(self.h)

    def results(self):
        n = self.num_features
        k = np.zeros((n, 3), np.float32)
        f = np.zeros((n, 4), np.float32)
        d = np.zeros((n, 128), np.float16)
        p = np.zeros((n, 128), np.float16)
        total = self._results(self.h, k.ctypes.data, f.ctypes.data, d.ctypes.data, p.ctypes.data)
        return dict(n=total, kpts=k, feats=f, desc=d, prev_desc=p)

    # --- stage level
    def stack(self, o: int, which: str) -> np.ndarray:
        layers = self.n_layers if which == "gauss" else self.n_layers - 1
        oc = self.octaves[o]
        out = np.zeros((layers, oc["h"], oc["w"]), np.float32)
        rc = self._get_stack(self.h, o, 0 if which == "gauss" else 1, out.ctypes.data)
        assert rc == 0
        return out

    def kpt_buffer(self, o: int, which: str) -> np.ndarray:
        cap = self.octaves[o]["cap"]
        spec = {"cand": (0, (cap, 3), np.float32), "candF": (1, (cap, 4), np.float32),
                "coll": (2, (cap, 3), np.float32), "collF": (3, (cap, 4), np.float32),
                "collMask": (4, (cap,), np.int32), "desc": (5, (cap, 128), np.float16)}[which]
        out = np.zeros(spec[1], spec[2])
        n = cap * 128 if which == "desc" else cap
        self._get_kpt_buffer(self.h, o, spec[0], out.ctypes.data, n)
        return out

    def taps(self, which: int) -> np.ndarray:
        out = np.zeros(128, np.float32)
        n = self._get_taps(self.h, which, out.ctypes.data, 128)
        return out[:n].copy()

    def match(self, des: np.ndarray, src: np.ndarray) -> np.ndarray:
        des = np.ascontiguousarray(des, dtype=np.float16)
        src = np.ascontiguousarray(src, dtype=np.float16)
        out = np.full(des.shape[0], -1, np.int32)
        self._match(des.ctypes.data, des.shape[0], src.ctypes.data, src.shape[0], out.ctypes.data)
        return out


class CpuOracle(_Oracle):
    prefix = "oracle_"

    def __init__(self, cols, rows, num_features=5000, remove_duplicates=False, interpolate_orientation=False,
                 working_upscale=False, **cfg):
        self.lib_path = build_cpu()
        self._opts = int(remove_duplicates) | (int(interpolate_orientation) << 1) | (int(working_upscale) << 2)
        super().__init__(cols, rows, num_features, **cfg)

    def counts(self, o: int):
        c = (C.c_int * 3)()
        self._get_counts(self.h, o, c)
        return dict(P=c[0], M=c[1], T=c[2])

    def set_threads(self, n: int):
        self.lib.oracle_set_threads(C.c_int(n))

    def threads(self) -> int:
        return int(self.lib.oracle_get_threads())


def _zero_free_device_memory():
    """The reference never zeroes its candidate / collected-keypoint buffers (plain cudaMalloc) and CONSUMES their stale
    last slots. In a fresh process cudaMalloc hands out zeroed memory; inside a long-lived process
    (a test suite) it hands out whatever earlier allocations left behind, and the reference then reads gradients at
    garbage coordinates: at best other results, at worst an illegal address that poisons the CUDA context for everything
    that follows. So before the reference allocates, grab (nearly) all free device memory, zero it and give it back: the
    reference starts from the state it has in a process of its own -- and from the state the engine, which zeroes its
    buffers, always starts from. Test infrastructure only."""
    try:
        rt = C.CDLL("libcudart.so")
    except OSError:
        try:
            rt = C.CDLL("libcudart.so.12")
        except OSError:
            return
    free, total = C.c_size_t(0), C.c_size_t(0)
    if rt.cudaMemGetInfo(C.byref(free), C.byref(total)) != 0:
        return
    want = int(free.value) - (1 << 30)
    while want > (1 << 28):
        p = C.c_void_p(0)
        if rt.cudaMalloc(C.byref(p), C.c_size_t(want)) == 0:
            rt.cudaMemset(p, 0, C.c_size_t(want))
            rt.cudaDeviceSynchronize()
            rt.cudaFree(p)
            return
        rt.cudaGetLastError()
        want //= 2


class RefOracle(_Oracle):
    prefix = "ref_"
    lib_path = REF_LIB

    def __init__(self, *args, **kwargs):
        _zero_free_device_memory()
        super().__init__(*args, **kwargs)

    def counts_raw(self, o: int):
        """(prefix_sum[last], prefix_sum[cap-1]); meaning depends on the stage just run."""
        c = (C.c_int * 2)()
        self._get_counts(self.h, o, c)
        return c[0], c[1]

    def time_device_resident(self, warmup: int, iters: int) -> float:
        fn = self.lib.ref_time_device_resident
        fn.restype, fn.argtypes = C.c_float, [C.c_void_p, C.c_int, C.c_int]
        return float(fn(self.h, warmup, iters))

    def time_e2e(self, warmup: int, iters: int, copy_to_host: bool = True) -> float:
        fn = self.lib.ref_time_e2e
        fn.restype, fn.argtypes = C.c_float, [C.c_void_p, C.c_int, C.c_int, C.c_int]
        return float(fn(self.h, warmup, iters, int(copy_to_host)))

    def time_match(self, des, src, warmup: int, iters: int) -> float:
        fn = self.lib.ref_time_match
        fn.restype = C.c_float
        fn.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int]
        des = np.ascontiguousarray(des, dtype=np.float16)
        src = np.ascontiguousarray(src, dtype=np.float16)
        return float(fn(des.ctypes.data, des.shape[0], src.ctypes.data, src.shape[0], warmup, iters))
