// TEST INFRASTRUCTURE -- not part of the product, never linked into it.
//
// C-ABI driver around the UNMODIFIED reference implementation
// (/root/reference/sift_cuda/...), built by oracle/Makefile into
// oracle/_ref/libref_sift.so. It is the strongest oracle we have for the hot
// path: the reference's own CUDA kernels compiled for sm_100 by this image's
// nvcc, run on the same inputs as the B200-native engine.
//
// The reference's Detector.cu is textually included so that the (inline, hence
// not exported) stage methods Detector::createInitialImage ... can be driven
// one by one and the private SiftData inspected between stages. No reference
// source is copied into this repository: the #include resolves into
// /root/reference at build time only.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <thrust/device_vector.h>
#include <thrust/host_vector.h>
#include <cub/device/device_scan.cuh>
#include <zlib.h>
#include <msgpack.hpp>

#define private public
#include "sift_cuda/interface/Detector.cu"
#undef private
#include "sift_cuda/sift_func/Match.cuh"

namespace {
struct RefHandle {
    std::unique_ptr<sift_cuda::Detector> det;
    Imagef img;
    std::vector<Imagef> frames;  // pre-built host images for the bench loop (the caller's Imagef objects)
    int rows = 0, cols = 0;
    int prev_size = 0;
};

template <typename T>
int copy_dev(const thrust::device_vector<T>& v, void* out, size_t max_elems) {
    size_t n = std::min(max_elems, v.size());
    if (out && n) {
        cudaMemcpy(out, thrust::raw_pointer_cast(v.data()), n * sizeof(T), cudaMemcpyDeviceToHost);
    }
    return int(v.size());
}
}  // namespace

extern "C" {

static void* ref_create_from(const CudaSiftConfig& cfg);

void* ref_create(int cols, int rows, int num_features) {
    CudaSiftConfig cfg;
    cfg.col_width = cols;
    cfg.row_width = rows;
    cfg.numFeatures = num_features;
    cfg.upscale = false;
    return ref_create_from(cfg);
}

// every field of CudaSiftConfig (types/CudaSiftConfig.hh:3-14) except upscale: the reference is parametric in
// numOctaveLayers and sigma (Detector.cu:63-94), these entry points let the tests drive it there
void* ref_create_cfg(int cols, int rows, int num_features, int num_octave_layers, double contrast, double edge, double sigma) {
    CudaSiftConfig cfg;
    cfg.col_width = cols;
    cfg.row_width = rows;
    cfg.numFeatures = num_features;
    cfg.numOctaveLayers = num_octave_layers;
    cfg.contrastThreshould = contrast;
    cfg.edgeThreshould = edge;
    cfg.sigma = sigma;
    cfg.upscale = false;
    return ref_create_from(cfg);
}

int ref_num_layers(void* p) { return static_cast<RefHandle*>(p)->det->m_config.numOctaveLayers + 3; }

static void* ref_create_from(const CudaSiftConfig& cfg) {
    const int rows = cfg.row_width, cols = cfg.col_width;
    auto* h = new RefHandle;
    h->rows = rows;
    h->cols = cols;
    // The constructor prints one line to stdout (Detector.hh:28); keep it off
    // the caller's stdout (bench.py prints exactly one JSON line there).
    std::streambuf* old = std::cout.rdbuf();
    std::ostringstream sink;
    std::cout.rdbuf(sink.rdbuf());
    h->det = std::make_unique<sift_cuda::Detector>(cfg);
    std::cout.rdbuf(old);
    if (!h->det->gpuWarmUpAndAllocate()) {
        delete h;
        return nullptr;
    }
    h->img.m_image_size = Size{float(rows), float(cols)};
    h->img.m_data = std::make_shared<std::vector<float>>(size_t(rows) * cols);
    cudaDeviceSynchronize();
    return h;
}

void ref_destroy(void* p) { delete static_cast<RefHandle*>(p); }

int ref_num_octaves(void* p) { return static_cast<RefHandle*>(p)->det->m_nOctaves; }

// dims[0..3] = width, height, pitch in floats (Gaussian stack), cap (kpt slots)
void ref_octave_info(void* p, int o, int* dims) {
    auto& d = *static_cast<RefHandle*>(p)->det;
    auto& g = d.m_data.gaussian_block_pyramid.at(o);
    dims[0] = int(g.getDim().x);
    dims[1] = int(g.getDim().y);
    dims[2] = int(g.getPitch() / sizeof(float));
    dims[3] = int(d.m_data.extremas.at(o).collectedKpts.size());
}

void ref_set_image(void* p, const float* img) {
    auto* h = static_cast<RefHandle*>(p);
    std::memcpy(h->img.m_data->data(), img, size_t(h->rows) * h->cols * sizeof(float));
}

// The reference's public entry point, unmodified: H2D + all stages.
int ref_detect(void* p) {
    auto* h = static_cast<RefHandle*>(p);
    h->det->detectAndCompute(h->img);
    cudaDeviceSynchronize();
    return h->det->total_size;
}

// Stage-by-stage replay of Detector::detectAndCompute (Detector.cu:133-233)
// with the same calls in the same order; `stage` selects how far to go:
// 0 upload+prev copy+initial blur, 1 gaussian, 2 dog, 3 extrema, 4 refine,
// 5 orientation, 6 descriptor, 7 gather. Each call runs exactly one stage.
int ref_stage(void* p, int stage) {
    auto* h = static_cast<RefHandle*>(p);
    auto& d = *h->det;
    switch (stage) {
        case 0:
            d.m_data.original_image = h->img;
            cudaMemcpyAsync(thrust::raw_pointer_cast(d.prev_descriptor.data()),
                            thrust::raw_pointer_cast(d.device_descriptor.data()),
                            d.prev_descriptor.size() * sizeof(half), cudaMemcpyDeviceToDevice);
            d.createInitialImage();
            break;
        case 1: d.getGaussianPyramid(); break;
        case 2: d.getDogPyramid(); break;
        case 3: d.getKeyPoints(); break;
        case 4: d.adjustKpts(); break;
        case 5: d.adjustOrientationAndCleanup(); break;
        case 6: d.extractDescriptor(); break;
        case 7: d.collectAcrossOctaves(); break;
        default: return -1;
    }
    cudaDeviceSynchronize();
    return cudaGetLastError() == cudaSuccess ? 0 : -2;
}

// Device-resident timing: stages 0' (initial blur only, no H2D) .. 7, `iters`
// times, CUDA events on the legacy default stream around the host-driven loop
// (the reference blocks the host between stages, so wall clock == its latency).
float ref_time_device_resident(void* p, int warmup, int iters) {
    auto* h = static_cast<RefHandle*>(p);
    auto& d = *h->det;
    d.m_data.original_image = h->img;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float total = 0.f;
    for (int i = 0; i < warmup + iters; i++) {
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        cudaMemcpyAsync(thrust::raw_pointer_cast(d.prev_descriptor.data()),
                        thrust::raw_pointer_cast(d.device_descriptor.data()),
                        d.prev_descriptor.size() * sizeof(half), cudaMemcpyDeviceToDevice);
        d.createInitialImage();
        d.getGaussianPyramid();
        d.getDogPyramid();
        d.getKeyPoints();
        d.adjustKpts();
        d.adjustOrientationAndCleanup();
        d.extractDescriptor();
        d.collectAcrossOctaves();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (i >= warmup) total += ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return total / float(iters);
}

// End-to-end timing through the public API (pageable H2D inside), host clock.
float ref_time_e2e(void* p, int warmup, int iters, int copy_to_host) {
    auto* h = static_cast<RefHandle*>(p);
    auto& d = *h->det;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float total = 0.f;
    for (int i = 0; i < warmup + iters; i++) {
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        d.detectAndCompute(h->img);
        if (copy_to_host) d.copyToHost(true);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (i >= warmup) total += ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return total / float(iters);
}

// --- bench helpers: the loop of tool/extract_and_match_example.cc:68-101 through the public API only.
int ref_add_frame(void* p, const float* img) {
    auto* h = static_cast<RefHandle*>(p);
    Imagef f;
    f.m_image_size = Size{float(h->rows), float(h->cols)};
    f.m_data = std::make_shared<std::vector<float>>(img, img + size_t(h->rows) * h->cols);
    h->frames.push_back(f);
    return int(h->frames.size()) - 1;
}
// detectAndCompute(host image) [+ copyToHost(descriptor)] [+ matchBruteForce(prev, cur)]; returns total_size.
// n_matched (may be NULL) receives the number of accepted matches (forces use of the returned vector).
int ref_process_frame(void* p, int idx, int copy_desc, int do_match, int* n_matched) {
    auto* h = static_cast<RefHandle*>(p);
    auto& d = *h->det;
    d.detectAndCompute(h->frames.at(idx));
    d.copyToHost(copy_desc != 0);
    cudaDeviceSynchronize();
    const int cur = d.total_size;
    int acc = 0;
    if (do_match && h->prev_size > 0 && cur > 0) {
        std::vector<int> m = sift_cuda::matchBruteForce(d.prev_descriptor, h->prev_size, d.device_descriptor, cur);
        for (int v : m) acc += v >= 0;
    }
    if (n_matched) *n_matched = acc;
    h->prev_size = cur;
    return cur;
}
// Device-resident variant: the frame is already in original_image; stages as in ref_time_device_resident.
int ref_process_resident(void* p, int do_match, int* n_matched) {
    auto* h = static_cast<RefHandle*>(p);
    auto& d = *h->det;
    cudaMemcpyAsync(thrust::raw_pointer_cast(d.prev_descriptor.data()), thrust::raw_pointer_cast(d.device_descriptor.data()),
                    d.prev_descriptor.size() * sizeof(half), cudaMemcpyDeviceToDevice);
    d.createInitialImage();
    d.getGaussianPyramid();
    d.getDogPyramid();
    d.getKeyPoints();
    d.adjustKpts();
    d.adjustOrientationAndCleanup();
    d.extractDescriptor();
    d.collectAcrossOctaves();
    cudaDeviceSynchronize();
    const int cur = d.total_size;
    int acc = 0;
    if (do_match && h->prev_size > 0 && cur > 0) {
        std::vector<int> m = sift_cuda::matchBruteForce(d.prev_descriptor, h->prev_size, d.device_descriptor, cur);
        for (int v : m) acc += v >= 0;
    }
    if (n_matched) *n_matched = acc;
    h->prev_size = cur;
    return cur;
}
// One frame of the bench loop with a wall-clock split (every phase ends in cudaDeviceSynchronize, so wall == device):
// ms4 = {H2D upload of frame idx (CudaImage::operator=, CudaImage.cu:88-105), the stage sequence of detectAndCompute on
// the resident image, copyToHost(true), matchBruteForce(previous frame -> this frame)}. The stages are the reference's
// own methods in the order of Detector.cu:133-233.
int ref_process_timed(void* p, int idx, int do_match, double* ms4) {
    auto* h = static_cast<RefHandle*>(p);
    auto& d = *h->det;
    using clk = std::chrono::steady_clock;
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    cudaDeviceSynchronize();
    const auto t0 = clk::now();
    d.m_data.original_image = h->frames.at(idx);
    cudaDeviceSynchronize();
    const auto t1 = clk::now();
    cudaMemcpyAsync(thrust::raw_pointer_cast(d.prev_descriptor.data()), thrust::raw_pointer_cast(d.device_descriptor.data()),
                    d.prev_descriptor.size() * sizeof(half), cudaMemcpyDeviceToDevice);
    d.createInitialImage();
    d.getGaussianPyramid();
    d.getDogPyramid();
    d.getKeyPoints();
    d.adjustKpts();
    d.adjustOrientationAndCleanup();
    d.extractDescriptor();
    d.collectAcrossOctaves();
    cudaDeviceSynchronize();
    const auto t2 = clk::now();
    d.copyToHost(true);
    cudaDeviceSynchronize();
    const auto t3 = clk::now();
    const int cur = d.total_size;
    if (do_match && h->prev_size > 0 && cur > 0) {
        std::vector<int> m = sift_cuda::matchBruteForce(d.prev_descriptor, h->prev_size, d.device_descriptor, cur);
        (void)m;
    }
    const auto t4 = clk::now();
    h->prev_size = cur;
    if (ms4) { ms4[0] = ms(t0, t1); ms4[1] = ms(t1, t2); ms4[2] = ms(t2, t3); ms4[3] = ms(t3, t4); }
    return cur;
}
void ref_upload_frame(void* p, int idx) {
    auto* h = static_cast<RefHandle*>(p);
    h->det->m_data.original_image = h->frames.at(idx);
    cudaDeviceSynchronize();
}

int ref_total_size(void* p) { return static_cast<RefHandle*>(p)->det->total_size; }

// Final results (device vectors, first total_size entries valid).
int ref_results(void* p, float* kpts3, float* feats4, uint16_t* desc, uint16_t* prev_desc) {
    auto& d = *static_cast<RefHandle*>(p)->det;
    cudaDeviceSynchronize();
    copy_dev(d.device_kpts, kpts3, d.device_kpts.size());
    copy_dev(d.device_features, feats4, d.device_features.size());
    copy_dev(d.device_descriptor, desc, d.device_descriptor.size());
    copy_dev(d.prev_descriptor, prev_desc, d.prev_descriptor.size());
    return d.total_size;
}

// Dense (unpadded) copy of one layer stack: which = 0 gaussian (6 layers), 1 dog (5 layers).
int ref_get_stack(void* p, int o, int which, float* out) {
    auto& d = *static_cast<RefHandle*>(p)->det;
    auto& b = which == 0 ? d.m_data.gaussian_block_pyramid.at(o) : d.m_data.dog_pyramid.at(o);
    auto dim = b.getDim();
    cudaError_t e = cudaMemcpy2D(out, size_t(dim.x) * 4, b.getData(), b.getPitch(), size_t(dim.x) * 4,
                                 size_t(dim.y) * size_t(dim.z), cudaMemcpyDeviceToHost);
    return e == cudaSuccess ? 0 : -1;
}

// Raw pitched copy incl. padding (to see what the reference reads at x == w, quirk Q12).
int ref_get_stack_pitched(void* p, int o, int which, float* out) {
    auto& d = *static_cast<RefHandle*>(p)->det;
    auto& b = which == 0 ? d.m_data.gaussian_block_pyramid.at(o) : d.m_data.dog_pyramid.at(o);
    auto dim = b.getDim();
    cudaError_t e = cudaMemcpy(out, b.getData(), b.getPitch() * size_t(dim.y) * size_t(dim.z), cudaMemcpyDeviceToHost);
    return e == cudaSuccess ? 0 : -1;
}

// Keypoint buffers of one octave. which: 0 candidateKpts(float3) 1 candidateFeatures(float4)
// 2 collectedKpts 3 collectedFeatures 4 collectedMask(int) 5 descriptor(half)
// 6 prefix_sum[0:cap] (int). Returns element count of the buffer.
int ref_get_kpt_buffer(void* p, int o, int which, void* out, int max_elems) {
    auto& k = static_cast<RefHandle*>(p)->det->m_data.extremas.at(o);
    switch (which) {
        case 0: return copy_dev(k.candidateKpts, out, max_elems);
        case 1: return copy_dev(k.candidateFeatures, out, max_elems);
        case 2: return copy_dev(k.collectedKpts, out, max_elems);
        case 3: return copy_dev(k.collectedFeatures, out, max_elems);
        case 4: return copy_dev(k.collectedMask, out, max_elems);
        case 5: return copy_dev(k.descriptor, out, max_elems);
        case 6: return copy_dev(k.prefix_sum, out, max_elems);
        default: return -1;
    }
}

// counts[0] = prefix_sum[last] (extrema count P right after stage 3),
// counts[1] = prefix_sum[cap-1] (M after stage 4, T after stage 5).
void ref_get_counts(void* p, int o, int* counts) {
    auto& k = static_cast<RefHandle*>(p)->det->m_data.extremas.at(o);
    cudaMemcpy(&counts[0], thrust::raw_pointer_cast(k.prefix_sum.data()) + k.prefix_sum.size() - 1, sizeof(int),
               cudaMemcpyDeviceToHost);
    cudaMemcpy(&counts[1], thrust::raw_pointer_cast(k.prefix_sum.data()) + k.collectedMask.size() - 1, sizeof(int),
               cudaMemcpyDeviceToHost);
}

// Gaussian taps as the reference generated them: which = -1 initial blur, 0..5 layer taps.
int ref_get_taps(void* p, int which, float* out, int max_elems) {
    auto& d = *static_cast<RefHandle*>(p)->det;
    const sift_cuda::CudaImage& k = which < 0 ? d.m_data.initial_image_kernel_1d : d.m_data.gaussian_kernels_1d.at(which);
    int n = int(k.getDim().x);
    if (out && n <= max_elems) cudaMemcpy(out, k.getData(), n * sizeof(float), cudaMemcpyDeviceToHost);
    return n;
}

// sift_cuda::matchBruteForce (Match.cu:8), host buffers in / out.
int ref_match(const uint16_t* des, int num_des, const uint16_t* src, int num_src, int* out_idx) {
    thrust::device_vector<half> d(size_t(num_des) * 128), s(size_t(num_src) * 128);
    cudaMemcpy(thrust::raw_pointer_cast(d.data()), des, size_t(num_des) * 256, cudaMemcpyHostToDevice);
    cudaMemcpy(thrust::raw_pointer_cast(s.data()), src, size_t(num_src) * 256, cudaMemcpyHostToDevice);
    std::vector<int> r = sift_cuda::matchBruteForce(d, num_des, s, num_src);
    std::memcpy(out_idx, r.data(), r.size() * sizeof(int));
    return int(r.size());
}

// Same call timed with CUDA events (includes the reference's per-call
// allocation of the score matrix and its cudaDeviceSynchronize, Match.cu:16-28).
float ref_time_match(const uint16_t* des, int num_des, const uint16_t* src, int num_src, int warmup, int iters) {
    thrust::device_vector<half> d(size_t(num_des) * 128), s(size_t(num_src) * 128);
    cudaMemcpy(thrust::raw_pointer_cast(d.data()), des, size_t(num_des) * 256, cudaMemcpyHostToDevice);
    cudaMemcpy(thrust::raw_pointer_cast(s.data()), src, size_t(num_src) * 256, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float total = 0.f;
    for (int i = 0; i < warmup + iters; i++) {
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        std::vector<int> r = sift_cuda::matchBruteForce(d, num_des, s, num_src);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (i >= warmup) total += ms;
    }
    return total / float(iters);
}

}  // extern "C"
