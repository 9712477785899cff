// Mirror of sift_cuda/sift_func/Match.cuh:9-14: std::vector<int> matchBruteForce(des, num_des, src, num_src),
// result[i] = index into src or -1, blocking -- over b200sift_match_device (tcgen05 filter + exact re-rank).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "../../b200sift.h"
#include "../types/DeviceView.hh"

namespace sift_cuda {

// Raw-pointer / stream variant (what the sharded path uses): device pointers, asynchronous on `stream`.
inline int matchBruteForceAsync(const half* des, int num_des, const half* src, int num_src, int* d_out_idx,
                                cudaStream_t stream = nullptr) {
    return b200sift_match_device(reinterpret_cast<const uint16_t*>(des), num_des, reinterpret_cast<const uint16_t*>(src),
                                 num_src, d_out_idx, stream);
}

// Works with sift_cuda::DeviceView<half> (Detector's members) and with thrust::device_vector<half>:
// anything whose .data() converts to a raw device pointer through device_ptr().
template <class DesVec, class SrcVec>
inline std::vector<int> matchBruteForce(const DesVec& des, const int num_des, const SrcVec& src, const int num_src) {
    std::vector<int> out(num_des > 0 ? num_des : 0, -1);
    if (num_des <= 0) return out;
    int* d_out = nullptr;
    auto die = [](const char* what) {  // CUDA_CHECK semantics of the reference (CudaMemRAII.cuh:11-19)
        std::fprintf(stderr, "matchBruteForce: %s: %s\n", what, b200sift_last_error());
        std::exit(EXIT_FAILURE);
    };
    if (cudaMalloc(&d_out, sizeof(int) * size_t(num_des)) != cudaSuccess) die("cudaMalloc");
    if (matchBruteForceAsync(device_ptr(des), num_des, device_ptr(src), num_src, d_out, nullptr) != B200SIFT_OK) die("match");
    if (cudaMemcpy(out.data(), d_out, sizeof(int) * size_t(num_des), cudaMemcpyDeviceToHost) != cudaSuccess) die("cudaMemcpy");
    cudaFree(d_out);
    return out;
}

}  // namespace sift_cuda
