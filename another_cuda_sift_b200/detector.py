"""Python mirror of the reference's public API for the SIFT hot path, over the C ABI.

Names, argument meaning and error behaviour follow the reference so that tests read like its own
usage (`/root/reference/readme.md:66-76`, `tool/extract_and_match_example.cc:57-101`):

    cfg = CudaSiftConfig(col_width=w, row_width=h)
    det = Detector(cfg)
    det.gpuWarmUpAndAllocate()
    det.detectAndCompute(img)          # img: float32 [rows, cols], values 0..255
    det.copyToHost(descriptor=True)    # det.final_kpts / final_features / descriptors, det.total_size
    idx = matchBruteForce(det.prev_descriptor, prev_n, det.device_descriptor, det.total_size)

Everything runs on the GPU through libb200sift.so; numpy / torch only carry buffers.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib
from ._lib import B200SiftError, check


@dataclass
class CudaSiftConfig:
    """`struct CudaSiftConfig` (reference sift_cuda/types/CudaSiftConfig.hh:3-14), same names."""

    col_width: int = 0
    row_width: int = 0
    numFeatures: int = 5000
    numOctaveLayers: int = 3
    contrastThreshould: float = 0.04
    edgeThreshould: float = 10.0
    sigma: float = 1.6
    upscale: bool = False

    def to_c(self) -> _lib.Config:
        return _lib.Config(self.col_width, self.row_width, self.numFeatures, self.numOctaveLayers,
                           self.contrastThreshould, self.edgeThreshould, self.sigma, int(self.upscale))


class DeviceArray:
    """Borrowed device buffer (owned by a Detector) exposing __cuda_array_interface__, so that
    `torch.as_tensor(arr, device="cuda")` / cupy can view it without a copy."""

    def __init__(self, ptr: int, shape, typestr: str, owner):
        self.ptr, self.shape, self.typestr, self._owner = ptr, tuple(shape), typestr, owner
        self.__cuda_array_interface__ = {"shape": self.shape, "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


def _device_ptr(x) -> int:
    if isinstance(x, DeviceArray):
        return x.ptr
    if hasattr(x, "data_ptr"):  # torch tensor
        if not x.is_cuda or not x.is_contiguous():
            raise B200SiftError("device tensor must be a contiguous CUDA tensor")
        return x.data_ptr()
    if hasattr(x, "__cuda_array_interface__"):
        return x.__cuda_array_interface__["data"][0]
    raise B200SiftError(f"not a device buffer: {type(x)}")


class Detector:
    """`sift_cuda::Detector` (reference sift_cuda/interface/Detector.hh:24-96)."""

    def __init__(self, config: CudaSiftConfig, *, frames_in_flight: int = 0, remove_duplicates: bool = False,
                 interpolate_orientation: bool = False, working_upscale: bool = False):
        """`config` is the reference's struct; the keyword arguments are b200sift_options (include/b200sift.h): engine
        options without a reference counterpart, all off = the reference's behaviour."""
        self.lib = _lib.load()
        self.config = config
        self.options = _lib.Options(frames_in_flight, int(remove_duplicates), int(interpolate_orientation), int(working_upscale))
        self._h = None
        self.total_size = 0
        self.max_kpts = 5000
        n = max(config.numFeatures, 1)
        # host mirrors (Detector.hh:58-60), sized numFeatures like the reference (Detector.cu:128-130)
        self.final_kpts = np.zeros((n, 3), np.float32)
        self.final_features = np.zeros((n, 4), np.float32)
        self.descriptors = np.zeros((n, 128), np.float16)
        self.device_kpts = self.device_features = self.device_descriptor = self.prev_descriptor = None
        self._dump_dir = None

    # -- lifecycle
    def gpuWarmUpAndAllocate(self) -> bool:
        """Detector.cu:17-39. Returns False (like the reference) when the image size is not set."""
        if self._h is not None:
            return True
        if self.config.col_width == 0 or self.config.row_width == 0:
            return False
        h = C.c_void_p()
        cfg = self.config.to_c()
        check(self.lib.b200sift_create_ex(C.byref(cfg), C.byref(self.options), C.byref(h)))
        self._h = h
        self.n_octaves = self.lib.b200sift_num_octaves(h)
        self.octaves = []
        for o in range(self.n_octaves):
            d = (C.c_int * 4)()
            check(self.lib.b200sift_octave_info(h, o, d))
            self.octaves.append(dict(w=d[0], h=d[1], pitch=d[2], cap=d[3]))
        self._refresh_views()
        return True

    def _refresh_views(self):
        """The engine keeps several frames in flight, each with its own result buffers: the public device views
        always point at the latest frame (prev_descriptor at the one before it, in place, no copy)."""
        k, f, d, p = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self.lib.b200sift_results(self._h, C.byref(k), C.byref(f), C.byref(d), C.byref(p), None))
        n = self.config.numFeatures
        self.device_kpts = DeviceArray(k.value, (n, 3), "<f4", self)
        self.device_features = DeviceArray(f.value, (n, 4), "<f4", self)
        self.device_descriptor = DeviceArray(d.value, (n, 128), "<f2", self)
        self.prev_descriptor = DeviceArray(p.value, (n, 128), "<f2", self)

    def close(self):
        if self._h is not None:
            self.lib.b200sift_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _handle(self):
        if self._h is None:
            raise B200SiftError("gpuWarmUpAndAllocate() has not been called")
        return self._h

    # -- hot path
    def detectAndCompute(self, image: np.ndarray) -> None:
        """Detector.cu:133-233. `image`: float32 (or uint8) [rows, cols]; blocks until done and sets
        total_size, like the reference (which synchronises six times inside this call)."""
        self.detectAndComputeAsync(image)
        self.synchronize()
        if self._dump_dir:
            self.dumpFrame(self._dump_dir)

    def setDataGen(self, path: str) -> None:
        """Detector.hh:48 / Detector.cu:145-229: once set, every blocking detectAndCompute writes the frame's
        intermediate stages under `path` (raw arrays + manifest.json instead of the reference's msgpack+zlib;
        tools/stage_replay.py is the counterpart of the reference's tool/perf.cu)."""
        self._dump_dir = path or None

    def dumpFrame(self, path: str) -> None:
        import os

        os.makedirs(path, exist_ok=True)
        check(self.lib.b200sift_dump_frame(self._handle(), os.fsencode(path)))

    def detectAndComputeAsync(self, image) -> None:
        h = self._handle()
        if hasattr(image, "is_cuda") and image.is_cuda:
            if image.dim() != 2 or image.dtype.is_floating_point is False or image.element_size() != 4:
                raise B200SiftError("device image must be float32 [rows, cols]")
            if image.stride(1) != 1:
                raise B200SiftError("device image rows must be contiguous")
            import torch

            # the engine orders its copy after the stream that produced the tensor (torch's current stream)
            check(self.lib.b200sift_detect_device_async(h, image.data_ptr(), image.stride(0) * 4,
                                                        C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            self._hold(image)  # the reference keeps the caching allocator off the tensor until its copy has long run
            return
        if hasattr(image, "numpy") and not isinstance(image, np.ndarray):
            image = image.numpy()  # CPU torch tensor (pinned or not) shares memory
        rows, cols = image.shape
        if image.dtype == np.uint8:
            image = np.ascontiguousarray(image)
            check(self.lib.b200sift_detect_host_u8_async(h, image.ctypes.data, rows, cols))
        else:
            image = np.ascontiguousarray(image, dtype=np.float32)
            check(self.lib.b200sift_detect_host_async(h, image.ctypes.data, rows, cols))
        self._hold(image)

    def _hold(self, image):
        self._keep = (getattr(self, "_keep", None) or [])[-self.frames_in_flight:] + [image]  # inputs must outlive their asynchronous copies
        self._refresh_views()

    def synchronize(self) -> int:
        t = C.c_int(0)
        check(self.lib.b200sift_sync(self._handle(), C.byref(t)))
        self.total_size = t.value
        return t.value

    def copyToHost(self, descriptor: bool = True) -> None:
        """Detector.cu:606-634: first total_size entries into final_kpts / final_features / descriptors."""
        self.synchronize()
        check(self.lib.b200sift_copy_to_host(self._handle(), self.final_kpts.ctypes.data,
                                             self.final_features.ctypes.data, self.descriptors.ctypes.data,
                                             int(descriptor)))

    # -- pipelined (batched-frame) use: nothing below blocks the host
    def matchConsecutiveAsync(self, out=None):
        """matchBruteForce(prev_descriptor, previous total, device_descriptor, this total) as
        tool/extract_and_match_example.cc:82 calls it, enqueued behind the current frame with both sizes
        read on the device. Returns a torch int32 CUDA tensor [numFeatures]; entries beyond the previous
        frame's total are -1. The tensor is written on the engine's stream (Detector.stream): order your own reads
        after it (synchronize(), or wait on that stream)."""
        import torch

        if out is None:
            out = torch.empty((self.config.numFeatures,), dtype=torch.int32, device="cuda")
        check(self.lib.b200sift_match_consecutive_async(self._handle(), out.data_ptr()))
        # the matcher writes `out` on the engine's stream, which torch's caching allocator knows nothing about: keep a
        # reference for as many frames as the engine has in flight, so the memory cannot be recycled under the kernel
        self._keep_out = (getattr(self, "_keep_out", None) or [])[-self.frames_in_flight:] + [out]
        return out

    def copyResultsAsync(self, host: "HostResults") -> None:
        """All numFeatures result slots + {total, previous total} to pinned host buffers, behind the frame."""
        check(self.lib.b200sift_copy_results_async(self._handle(), host.kpts.data_ptr(), host.feats.data_ptr(),
                                                   host.desc.data_ptr(), host.counts.data_ptr()))

    def detectAndComputeBatch(self, frames, results, match_out=None):
        """Frames of a batch, pipelined, through ONE native call (b200sift_detect_batch_host: the loop of
        tool/extract_and_match_example.cc:57-101 without a host synchronisation per frame): the host->device copy
        of frame i+1 overlaps the graph of frame i; results[i] (HostResults) receives frame i; match_out[i] (pinned
        int32 torch tensors), when given, receives the matches of frame i-1 -> frame i. `frames`: float32 or uint8
        [rows, cols] numpy arrays / CPU torch tensors (pinned for the copies to overlap), all of one dtype."""
        n = len(frames)
        if n == 0:
            return
        arrs = [f.numpy() if (hasattr(f, "numpy") and not isinstance(f, np.ndarray)) else f for f in frames]
        is_u8 = arrs[0].dtype == np.uint8
        arrs = [np.ascontiguousarray(a, dtype=np.uint8 if is_u8 else np.float32) for a in arrs]
        rows, cols = arrs[0].shape
        if any(a.shape != (rows, cols) for a in arrs):
            raise B200SiftError("all frames of a batch must have the same size")
        ptrs = lambda vals: (C.c_void_p * n)(*vals)
        check(self.lib.b200sift_detect_batch_host(
            self._handle(), ptrs([a.ctypes.data for a in arrs]), n, rows, cols, int(is_u8),
            ptrs([r.kpts.data_ptr() for r in results]), ptrs([r.feats.data_ptr() for r in results]),
            ptrs([r.desc.data_ptr() for r in results]), ptrs([r.counts.data_ptr() for r in results]),
            ptrs([m.data_ptr() for m in match_out]) if match_out is not None else None))
        self._keep = arrs[-self.frames_in_flight:]
        self.total_size = int(results[-1].counts[0])
        self._refresh_views()

    @property
    def frames_in_flight(self) -> int:
        if getattr(self, "_fif", None) is None:
            self._fif = int(self.lib.b200sift_frames_in_flight(self._h))
        return self._fif

    @property
    def stream(self) -> int:
        return int(self.lib.b200sift_stream(self._handle()) or 0)

    # -- introspection used by tests / bench (no reference counterpart besides setDataGen dumps)
    def launches_per_frame(self) -> int:
        return check(self.lib.b200sift_launches_per_frame(self._handle()))

    def time_frames(self, warmup: int, iters: int) -> float:
        """Latency of one frame graph (ms): the latest frame's graph replayed back to back."""
        ms = C.c_float(0)
        check(self.lib.b200sift_time_frames(self._handle(), warmup, iters, C.byref(ms)))
        return float(ms.value)

    def time_frames_pipelined(self, warmup: int, iters: int) -> float:
        """ms per frame with every frame slot in flight (every slot replays its graph, round robin)."""
        ms = C.c_float(0)
        check(self.lib.b200sift_time_frames_pipelined(self._handle(), warmup, iters, C.byref(ms)))
        return float(ms.value)

    KERNEL_CLASSES = ("copy", "blur_oct0", "blur_small", "downsample", "detect", "compact", "refine", "orient", "describe")

    def profile_frame(self, warmup: int, iters: int) -> dict:
        """ms per kernel class of one frame (CUDA events after every launch, eager frame)."""
        n = len(self.KERNEL_CLASSES)
        ms, cnt = (C.c_float * n)(), (C.c_int * n)()
        check(self.lib.b200sift_profile_frame(self._handle(), warmup, iters, ms, cnt))
        return {k: dict(ms=float(ms[i]), launches=int(cnt[i])) for i, k in enumerate(self.KERNEL_CLASSES)}

    def time_stages(self, iters: int = 50) -> dict:
        """Cumulative in-graph ms after pyramid / extrema / refine / orientation / descriptor."""
        ms = (C.c_float * 5)()
        check(self.lib.b200sift_time_stages(self._handle(), iters, ms))
        return dict(zip(("pyramid", "extrema", "refine", "orientation", "descriptor"), [float(v) for v in ms]))

    def debug_run_stages(self, image: np.ndarray, last_stage: int) -> None:
        image = np.ascontiguousarray(image, dtype=np.float32)
        check(self.lib.b200sift_debug_run_stages(self._handle(), image.ctypes.data, image.shape[0], image.shape[1], last_stage))
        self._refresh_views()

    def debug_gauss(self, o: int) -> np.ndarray:
        oc = self.octaves[o]
        out = np.zeros((self.config.numOctaveLayers + 3, oc["h"], oc["w"]), np.float32)
        check(self.lib.b200sift_debug_get_stack(self._handle(), o, 0, out.ctypes.data))
        return out

    def debug_kpt_buffer(self, o: int, which: str) -> np.ndarray:
        cap = self.octaves[o]["cap"]
        idx, width = {"cand": (0, 3), "candF": (1, 4), "coll": (2, 3), "collF": (3, 4)}[which]
        out = np.zeros((cap, width), np.float32)
        check(self.lib.b200sift_debug_get_kpt_buffer(self._handle(), o, idx, out.ctypes.data, cap))
        return out

    def debug_counts(self, o: int) -> dict:
        c = (C.c_int * 3)()
        check(self.lib.b200sift_debug_get_counts(self._handle(), o, c))
        return dict(P=c[0], M=c[1], T=c[2])

    def debug_taps(self, which: int) -> np.ndarray:
        out = np.zeros(128, np.float32)
        n = check(self.lib.b200sift_debug_get_taps(self._handle(), which, out.ctypes.data, 128))
        return out[:n].copy()


class HostResults:
    """Pinned host buffers for one frame's results (Detector.copyResultsAsync): kpts [n,3] f32, feats [n,4] f32,
    desc [n,128] f16, counts [2] i32 = {total, previous frame's total}. Only the first counts[0] rows are valid."""

    def __init__(self, num_features: int):
        import torch

        self.kpts = torch.zeros(num_features, 3, dtype=torch.float32).pin_memory()
        self.feats = torch.zeros(num_features, 4, dtype=torch.float32).pin_memory()
        self.desc = torch.zeros(num_features, 128, dtype=torch.float16).pin_memory()
        self.counts = torch.zeros(2, dtype=torch.int32).pin_memory()

    @property
    def total(self) -> int:
        return int(self.counts[0])

    def nbytes(self) -> int:
        return self.kpts.numel() * 4 + self.feats.numel() * 4 + self.desc.numel() * 2 + 8


def matchBruteForce(des, num_des: int, src, num_src: int, stream: int = 0, out=None):
    """`sift_cuda::matchBruteForce` (reference sift_func/Match.cuh:9-14).

    Host inputs (numpy float16 [n,128]) -> blocking, returns numpy int32 [num_des] like the
    reference's std::vector<int>. Device inputs (torch CUDA tensors / DeviceArray) -> enqueued on
    `stream`; returns `out` (a torch int32 CUDA tensor, allocated if not given) without blocking.
    """
    lib = _lib.load()
    if isinstance(des, np.ndarray):
        des = np.ascontiguousarray(des[:num_des], dtype=np.float16)
        src = np.ascontiguousarray(src[:num_src], dtype=np.float16)
        res = np.full(num_des, -1, np.int32)
        check(lib.b200sift_match_host(des.ctypes.data, num_des, src.ctypes.data, num_src, res.ctypes.data))
        return res
    import torch

    if out is None:
        out = torch.empty(max(num_des, 1), dtype=torch.int32, device="cuda")
    check(lib.b200sift_match_device(_device_ptr(des), num_des, _device_ptr(src), num_src, out.data_ptr(),
                                    C.c_void_p(stream)))
    return out[:num_des]
