// Mirror of the reference's sift_cuda::Detector (sift_cuda/interface/Detector.hh:24-96): same class name,
// constructor, gpuWarmUpAndAllocate / detectAndCompute / copyToHost / setDataGen, and the same public result
// members, implemented header-only over the C ABI (include/b200sift.h). Code written against the reference
// (readme.md:66-76, tool/detection_example.cc:39-49, tool/extract_and_match_example.cc:57-101) compiles against
// this header by changing the include path and linking libb200sift.so instead of the reference's objects.
//
// Differences a caller can observe:
//  * device result members are sift_cuda::DeviceView<T> (borrowed, owned by the engine) instead of
//    thrust::device_vector<T>; thrust::raw_pointer_cast is not needed, use .data();
//  * detectAndCompute returns with the frame finished (the reference returns with its last kernel in flight);
//  * setDataGen writes raw arrays + manifest.json (b200sift_dump_frame) instead of msgpack+zlib files;
//  * the reference's public stage methods (createInitialImage ... collectAcrossOctaves) do not exist: the
//    frame is one CUDA graph.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <thrust/host_vector.h>

#include "../../b200sift.h"
#include "../sift_func/Match.cuh"
#include "../types/CudaSiftConfig.hh"
#include "../types/DeviceView.hh"
#include "../types/HostImage.hh"

namespace sift_cuda {

class Detector {
  public:
    Detector(const CudaSiftConfig& config) : m_config(config) {
        // Detector.hh:27-28 (the reference prints this line too)
        m_nOctaves = int(std::round(std::log(double(std::min(m_config.col_width * 2, m_config.row_width * 2))) / std::log(2.) - 2) + 1);
        std::cout << "nOctaves: " << m_nOctaves << ". " << m_config.col_width << ", " << m_config.row_width << std::endl;
    }
    ~Detector() {
        if (m_handle) b200sift_destroy(m_handle);
    }
    Detector(const Detector&) = delete;
    Detector& operator=(const Detector&) = delete;

    // Detector.cu:17-39: false when the image size is not set, true otherwise; idempotent.
    bool gpuWarmUpAndAllocate() {
        if (m_initialized) return true;
        if (m_config.col_width == 0 || m_config.row_width == 0) {
            std::cerr << "Image width or height not set." << std::endl;  // Detector.cu:23
            return false;
        }
        b200sift_config c;
        b200sift_default_config(&c);
        c.col_width = m_config.col_width;
        c.row_width = m_config.row_width;
        c.numFeatures = m_config.numFeatures;
        c.numOctaveLayers = m_config.numOctaveLayers;
        c.contrastThreshould = m_config.contrastThreshould;
        c.edgeThreshould = m_config.edgeThreshould;
        c.sigma = m_config.sigma;
        c.upscale = m_config.upscale ? 1 : 0;
        check(b200sift_create(&c, &m_handle), "b200sift_create");
        refresh_views();
        const size_t n = size_t(m_config.numFeatures);
        final_kpts.resize(n);        // Detector.cu:128-130
        final_features.resize(n);
        descriptors.resize(n * 128);
        m_initialized = true;
        return true;
    }

    // The engine keeps several frames in flight (b200sift_frames_in_flight), each with its own result buffers: after every detect call the
    // public views point at the latest frame, prev_descriptor at the frame before it (in place, no copy).
    void refresh_views() {
        const float *k = nullptr, *f = nullptr;
        const uint16_t *d = nullptr, *p = nullptr;
        check(b200sift_results(m_handle, &k, &f, &d, &p, nullptr), "b200sift_results");
        const size_t n = size_t(m_config.numFeatures);
        device_kpts = DeviceView<float3>(reinterpret_cast<const float3*>(k), n);
        device_features = DeviceView<float4>(reinterpret_cast<const float4*>(f), n);
        device_descriptor = DeviceView<half>(reinterpret_cast<const half*>(d), n * 128);
        prev_descriptor = DeviceView<half>(reinterpret_cast<const half*>(p), n * 128);
    }

    // Detector.cu:133-233. The image must have the configured size (checked here, unchecked in the reference).
    void detectAndCompute(const Imagef& image) {
        require_init();
        check(b200sift_detect_host(m_handle, image.m_data->data(), int(image.m_image_size.row), int(image.m_image_size.col),
                                   &total_size), "b200sift_detect_host");
        refresh_views();
        if (!m_debug_path.empty()) check(b200sift_dump_frame(m_handle, m_debug_path.c_str()), "b200sift_dump_frame");
    }
    // Extension: frame already resident in HBM (pitched fp32), produced on `producer` (NULL = legacy default stream);
    // the engine orders its copy after that stream and enqueues the frame on stream(); call synchronize().
    void detectAndComputeDevice(const float* d_image, size_t pitch_bytes, cudaStream_t producer = nullptr) {
        require_init();
        check(b200sift_detect_device_async(m_handle, d_image, pitch_bytes, producer), "b200sift_detect_device_async");
        refresh_views();
    }
    int synchronize() {
        require_init();
        check(b200sift_sync(m_handle, &total_size), "b200sift_sync");
        return total_size;
    }
    cudaStream_t stream() { return static_cast<cudaStream_t>(b200sift_stream(m_handle)); }

    // Detector.cu:606-634: first total_size entries of kpts / features (/ descriptors) to the host mirrors.
    void copyToHost(bool descriptor) {
        require_init();
        check(b200sift_copy_to_host(m_handle, reinterpret_cast<float*>(final_kpts.data()),
                                    reinterpret_cast<float*>(final_features.data()),
                                    reinterpret_cast<uint16_t*>(descriptors.data()), descriptor ? 1 : 0), "b200sift_copy_to_host");
    }

    // Detector.hh:48: once set, every detectAndCompute dumps its intermediate stages into the (existing) directory
    // `path`: raw arrays + manifest.json (b200sift_dump_frame) instead of the reference's msgpack+zlib files.
    void setDataGen(const std::string& path) { m_debug_path = path; }

    // Results (Detector.hh:54-62)
    DeviceView<float3> device_kpts;       // x, y in full-resolution pixels, z = layer
    DeviceView<float4> device_features;   // x = packed octave, y = size, z = |contrast|, w = angle (deg)
    DeviceView<half> device_descriptor;   // N x 128, values 0..512
    DeviceView<half> prev_descriptor;     // previous frame's descriptors (Detector.cu:136-141)
    thrust::host_vector<float3> final_kpts{};
    thrust::host_vector<float4> final_features{};
    thrust::host_vector<half> descriptors{};
    int max_kpts = 5000;
    int total_size{0};

  private:
    void require_init() {
        if (!m_initialized) {
            std::fprintf(stderr, "sift_cuda::Detector used before gpuWarmUpAndAllocate()\n");
            std::exit(EXIT_FAILURE);
        }
    }
    static void check(int rc, const char* what) {  // CUDA_CHECK semantics (CudaMemRAII.cuh:11-19): print + exit
        if (rc != B200SIFT_OK) {
            std::fprintf(stderr, "%s failed (%d): %s\n", what, rc, b200sift_last_error());
            std::exit(EXIT_FAILURE);
        }
    }
    CudaSiftConfig m_config{};
    bool m_initialized{false};
    int m_nOctaves{0};
    std::string m_debug_path;
    b200sift_detector* m_handle{nullptr};
};

}  // namespace sift_cuda
