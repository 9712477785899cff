// b200sift_match_device_noalloc: the brute-force matcher with its scratch memory in a file-scope device array, for
// callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// No counterpart in the reference (its matcher allocates a thrust::device_vector per call, sift_func/Match.cu:16-28).
// It is the SAME matcher as b200sift_match_device (match_launch in
// match_tc.cu: norms -> tcgen05 filter -> exact re-rank, bit-exact with the reference); only where the survivor records
// and the squared norms live differs: b200sift_match_device keeps a cudaMalloc'd buffer per device that grows on demand,
// this entry point uses g_noalloc_scratch below and therefore has a fixed maximum problem size.
//
// Rules that follow from the design:
//  * every GPU that loads the library holds the array (kNoAllocBytes, ~8.1 MiB) until the process exits;
//  * in HOST code the array's name is a host-side stand-in, not a device address: the device address is obtained once per
//    device, with that device current, by cudaGetSymbolAddress (noalloc_scratch() below) and passed to the kernels;
//  * every call on a device shares the array and the array keeps the last call's contents. Nothing in it has to start at
//    zero: a call writes the squared norms of all its rows and one survivor record per (query row, column chunk) before
//    its re-rank kernel reads them, and reads nothing else. Calls that overlap on the GPU (different streams, or the
//    same stream from two threads without ordering) race on it and return garbage: order them (one stream, or events);
//  * the call path contains no cudaMalloc / cudaFree / allocating library routine: cudaGetDevice, cudaGetSymbolAddress
//    (first call per device), cudaFuncSetAttribute (first call per device and kernel) and three kernel launches.
#include <mutex>

#include "../../include/b200sift.h"
#include "common.cuh"

namespace b200sift {
int set_error(int code, const char* fmt, ...);  // engine.cu
}

namespace {

constexpr int kNoAllocMaxRows = B200SIFT_NOALLOC_MAX_ROWS;  // query rows and train rows, each
// survivor records (32 chunks x 32 bytes per query row) + squared norms of both sets, rounded up generously
constexpr size_t kNoAllocBytes = (size_t)kNoAllocMaxRows * 32 * 32 + 128 * 1024;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

constexpr int kMaxDev = 64;
std::mutex g_mu;
void* g_ptr[kMaxDev] = {nullptr};  // device address of g_noalloc_scratch on device i (each device has its own copy)

// Device address of the scratch array on the CURRENT device; resolved once per device.
int noalloc_scratch(void** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return b200sift::set_error(B200SIFT_ECUDA, "cudaGetDevice: %s (this engine has no CPU fallback)", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDev) return b200sift::set_error(B200SIFT_EINVAL, "device index out of range");
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return b200sift::set_error(B200SIFT_ECUDA, "cudaGetSymbolAddress(g_noalloc_scratch): %s", cudaGetErrorString(e));
        if (b200sift::match_scratch_bytes(kNoAllocMaxRows, kNoAllocMaxRows) > kNoAllocBytes)
            return b200sift::set_error(B200SIFT_ESTATE, "internal: the static matcher scratch is smaller than its stated maximum needs");
        g_ptr[dev] = p;
    }
    *out = g_ptr[dev];
    return 0;
}

}  // namespace

extern "C" {

int b200sift_match_noalloc_max_rows(void) { return kNoAllocMaxRows; }

int b200sift_match_device_noalloc(const uint16_t* d_query, int nq, const uint16_t* d_train, int nt, int* d_out_idx, void* stream) {
    if (nq < 0 || nt < 0 || (nq > 0 && (!d_query || !d_out_idx)) || (nt > 0 && !d_train))
        return b200sift::set_error(B200SIFT_EINVAL, "bad argument");
    if (nq > kNoAllocMaxRows || nt > kNoAllocMaxRows)
        return b200sift::set_error(B200SIFT_EINVAL,
                                   "b200sift_match_device_noalloc: %d x %d exceeds the static scratch (at most %d query rows and %d train rows); "
                                   "use b200sift_match_device for larger sets",
                                   nq, nt, kNoAllocMaxRows, kNoAllocMaxRows);
    if (nq == 0) return 0;
    void* scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;
    const size_t need = b200sift::match_scratch_bytes(nq, nt);
    if (need > kNoAllocBytes) return b200sift::set_error(B200SIFT_EINVAL, "b200sift_match_device_noalloc: problem needs %zu bytes of scratch, the static array holds %zu", need, kNoAllocBytes);
    const int rc = b200sift::match_launch(reinterpret_cast<const __half*>(d_query), nq, reinterpret_cast<const __half*>(d_train), nt, d_out_idx,
                                          scratch, kNoAllocBytes, static_cast<cudaStream_t>(stream));
    if (rc) return b200sift::set_error(B200SIFT_EINVAL, "b200sift_match_device_noalloc: matcher rejected the launch (%d)", rc);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return b200sift::set_error(B200SIFT_ECUDA, "b200sift_match_device_noalloc: %s", cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
