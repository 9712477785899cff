// Host side of the B200 SIFT engine: allocation, per-frame CUDA graph, C ABI (include/b200sift.h).
//
// Replaces the orchestration of interface/Detector.cu (5 CUDA graphs, 6 host synchronisations and
// a pinned-memory count round trip per frame) with ONE graph per frame and no host involvement
// between its first and last kernel: every count stays on the device.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200sift.h"
#include "common.cuh"

using namespace b200sift;

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                                  \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail(B200SIFT_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, cudaGetErrorName(e_), \
                        cudaGetErrorString(e_));                                                  \
    } while (0)

// get1DGaussian (utils/GaussianUtils.cc:39-67), float arithmetic on the host, bit-identical taps.
Taps gaussian_taps(float sigma) {
    Taps t;
    std::memset(&t, 0, sizeof(t));
    const int size = int(std::round(sigma * 6 + 1)) | 1;
    t.n = size;
    if (size > kMaxTaps) return t;
    float mean = size / 2;
    float cov_double = sigma * sigma * 2;
    float sum = 0;
    int i = 0;
    for (float row = 0; row < size; row++) {
        float g = expf(-((powf(row - mean, 2)) / cov_double));
        sum += g;
        t.v[i++] = g;
    }
    for (int k = 0; k < size; k++) t.v[k] = t.v[k] / sum;
    return t;
}

bool radius_supported(int n) {
    const int r = n / 2;
    return (n & 1) && (r == 4 || r == 5 || r == 6 || r == 8 || r == 10);
}

inline int pitch_floats(int w) { return ((w * 4 + 511) / 512) * 512 / 4; }

}  // namespace

constexpr int kSideStreams = 6;   // side branches of the frame graph
constexpr int kBranchPixels = 65536;  // an octave with at least this many pixels gets a branch of its own

struct b200sift_detector {
    b200sift_config cfg;
    int device = 0;
    int nOct = 0;
    Taps init_taps;
    Taps taps[kGaussLayers];
    FrameParams fp;
    // Two input buffers + two captured frame graphs: the H2D copy of frame i+1 runs on the copy stream while
    // the graph of frame i runs on the main stream.
    float* d_input[2] = {nullptr, nullptr};
    int in_pitch = 0;
    uint8_t* d_input_u8[2] = {nullptr, nullptr};
    size_t u8_pitch = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr};
    unsigned long long frame_no = 0;
    int* d_rowcount_all = nullptr;
    size_t rowcount_bytes = 0;
    __half* d_prev_desc = nullptr;
    int* h_total = nullptr;  // pinned
    void* match_scratch = nullptr;  // survivor records of the consecutive-frame matcher
    size_t match_scratch_bytes_ = 0;
    std::vector<void*> allocs;
    cudaStream_t stream = nullptr;
    cudaStream_t side[kSideStreams] = {nullptr};
    cudaEvent_t ev_fork[kMaxOctaves] = {nullptr}, ev_join[kMaxOctaves] = {nullptr}, ev_counts[kMaxOctaves] = {nullptr};
    int tail_first = 0;  // first octave handled by the single-CTA tail kernel
    // Branches of the frame graph: octave ranges [br_begin, br_end). Branches 0 .. n_side-1 cover the
    // multi-CTA octaves and run on side streams; the tail octaves (if any) run on the main stream.
    int n_side = 0;
    int br_begin[kSideStreams] = {0}, br_end[kSideStreams] = {0};
    cudaGraph_t graph[2] = {nullptr, nullptr};
    cudaGraphExec_t exec[2] = {nullptr, nullptr};
    int launches = 0;
    bool have_frame = false;
};

namespace {

// Optional per-launch event marks for b200sift_profile_frame: the time between two consecutive marks is
// charged to the kernel class of the later one.
struct Profiler {
    std::vector<cudaEvent_t> ev;
    std::vector<int> cls;
    size_t used = 0;
    cudaStream_t s = nullptr;
    void mark(int c) {
        if (used == ev.size()) {
            cudaEvent_t e;
            cudaEventCreate(&e);
            ev.push_back(e);
            cls.push_back(c);
        }
        cls[used] = c;
        cudaEventRecord(ev[used++], s);
    }
    ~Profiler() { for (auto e : ev) cudaEventDestroy(e); }
};
#define MARK(c) do { if (prof) prof->mark(c); } while (0)

template <typename T>
int dev_alloc(b200sift_detector* d, T** p, size_t count) {
    void* q = nullptr;
    const size_t bytes = std::max<size_t>(count * sizeof(T), 16);
    CU(cudaMalloc(&q, bytes));
    CU(cudaMemset(q, 0, bytes));
    d->allocs.push_back(q);
    *p = static_cast<T*>(q);
    return 0;
}

// The frame pipeline, stages numbered as in Detector::detectAndCompute (Detector.cu:133-233):
// 0 prev-descriptor copy, 1 pyramid, 3 extrema, 4 refinement, 5 orientation, 6/7 descriptor+gather.
int enqueue_frame(b200sift_detector* d, cudaStream_t s, int last_stage, int* launches, Profiler* prof = nullptr, int buf = 0) {
    FrameParams& fp = d->fp;
    int n = 0;
    MARK(B200SIFT_K_COPY);
    CU(cudaMemcpyAsync(fp.total + 1, fp.total, sizeof(int), cudaMemcpyDeviceToDevice, s));  // previous frame's count
    CU(cudaMemcpyAsync(d->d_prev_desc, fp.out_desc, (size_t)fp.numFeatures * 128 * sizeof(__half),
                       cudaMemcpyDeviceToDevice, s));
    CU(cudaMemsetAsync(d->d_rowcount_all, 0, d->rowcount_bytes, s));
    MARK(B200SIFT_K_COPY);
    // Gaussian pyramid. Octave o+1 only needs layer 3 of octave o, so layers 4 and 5 of every big octave
    // go to a side stream (a parallel branch of the frame graph) while the main chain moves on to the next
    // octave; octaves of <= kTailPixels pixels are walked by one CTA (launch_pyramid_tail).
    // Everything after the pyramid is per-octave work, and the small octaves are a long chain of
    // latency-bound launches. The frame graph therefore branches per octave range (build(): one branch
    // per big octave, one for the remaining multi-CTA octaves, the tail octaves on the main stream); each
    // branch runs detect .. orientation .. descriptor as soon as the six layers of its octaves exist.
    // A descriptor launch needs the keypoint counts of all lower octaves (its offset into the output), so
    // branch b waits for ev_counts[b-1], which branch b-1 records after its own wait: the dependency is
    // transitive. The branch that ends at the last octave publishes the frame's keypoint count.
    const bool fork = (prof == nullptr) && d->side[0] != nullptr;
    auto chain = [&](int ob, int oe, cudaStream_t cs) -> int {
        if (last_stage >= 3) {
            launch_detect(fp, ob, oe, cs); n++; MARK(B200SIFT_K_DETECT);
            launch_scatter(fp, ob, oe, cs); n++;
            MARK(B200SIFT_K_COMPACT);
        }
        if (last_stage >= 4) {
            launch_refine(fp, ob, oe, cs); n++;
            MARK(B200SIFT_K_REFINE);
            launch_compact_refined(fp, ob, oe, cs); n++;
            MARK(B200SIFT_K_COMPACT);
        }
        if (last_stage >= 5) {
            launch_orientation(fp, ob, oe, cs); n++;
            MARK(B200SIFT_K_ORIENT);
            launch_compact_oriented(fp, ob, oe, cs); n++;
            MARK(B200SIFT_K_COMPACT);
        }
        return 0;
    };
    int br = 0;  // branch of the current multi-CTA octave
    bool used[kSideStreams] = {false};
    for (int o = 0; o < d->nOct; o++) {
        OctParams& oc = fp.oct[o];
        const size_t ls = (size_t)oc.pitch * oc.h;
        const bool tail = o >= d->tail_first;
        if (o == 0) {
            launch_blur(d->d_input[buf], d->in_pitch, oc.gauss, oc.pitch, oc.w, oc.h, d->init_taps, s);
            MARK(B200SIFT_K_BLUR);
            n++;
        } else if (!tail || o == d->tail_first) {
            const OctParams& pv = fp.oct[o - 1];
            const float* src = pv.gauss + (size_t)pv.pitch * pv.h * kScales;
            launch_downsample(src, pv.pitch, pv.w, pv.h, oc.gauss, oc.pitch, oc.w, oc.h, s);
            MARK(B200SIFT_K_DOWNSAMPLE);
            n++;
        }
        if (tail) {
            if (o == d->tail_first) {
                launch_pyramid_tail(fp, d->taps, o, s);
                MARK(B200SIFT_K_BLUR_SMALL);
                n++;
            }
            continue;
        }
        while (o >= d->br_end[br]) br++;
        cudaStream_t bs = fork ? d->side[br] : s;  // the branch's stream
        cudaStream_t ls_stream = s;
        for (int l = 1; l < kGaussLayers; l++) {
            if (l == 4 && fork) {  // layers 4,5 off the critical path
                CU(cudaEventRecord(d->ev_fork[o], s));
                CU(cudaStreamWaitEvent(bs, d->ev_fork[o], 0));
                ls_stream = bs;
                used[br] = true;
            }
            launch_blur(oc.gauss + ls * (l - 1), oc.pitch, oc.gauss + ls * l, oc.pitch, oc.w, oc.h, d->taps[l], ls_stream);
            MARK(o == 0 ? B200SIFT_K_BLUR : B200SIFT_K_BLUR_SMALL);
            n++;
        }
        if (o == d->br_end[br] - 1) {  // the branch's octaves are complete
            if (int rc = chain(d->br_begin[br], d->br_end[br], bs)) return rc;
            if (fork) {
                if (br > 0) CU(cudaStreamWaitEvent(bs, d->ev_counts[br - 1], 0));
                CU(cudaEventRecord(d->ev_counts[br], bs));
            }
            if (last_stage >= 6) {
                CU(launch_descriptor(fp, d->br_begin[br], d->br_end[br], bs));
                n++;
                MARK(B200SIFT_K_DESCRIBE);
            }
        }
    }
    if (d->tail_first < d->nOct) {  // tail octaves on the main stream
        if (int rc = chain(d->tail_first, d->nOct, s)) return rc;
        if (fork && d->n_side > 0) CU(cudaStreamWaitEvent(s, d->ev_counts[d->n_side - 1], 0));
        if (last_stage >= 6) {
            CU(launch_descriptor(fp, d->tail_first, d->nOct, s));
            n++;
            MARK(B200SIFT_K_DESCRIBE);
        }
    }
    for (int b = 0; b < d->n_side; b++)
        if (used[b]) {
            CU(cudaEventRecord(d->ev_join[b], d->side[b]));
            CU(cudaStreamWaitEvent(s, d->ev_join[b], 0));
        }
    if (last_stage >= 6) {
        CU(cudaMemcpyAsync(d->h_total, fp.total, sizeof(int), cudaMemcpyDeviceToHost, s));
        MARK(B200SIFT_K_COPY);
    }
    CU(cudaGetLastError());
    if (launches) *launches = n;
    return 0;
}

int build(b200sift_detector* d) {
    const b200sift_config& c = d->cfg;
    CU(cudaGetDevice(&d->device));
    CU(cudaFree(0));
    // Detector.hh:27
    d->nOct = (int)(std::round(std::log(double(std::min(c.col_width * 2, c.row_width * 2))) / std::log(2.) - 2) + 1);
    if (d->nOct < 1 || d->nOct > kMaxOctaves) return fail(B200SIFT_EINVAL, "unsupported image size (%d octaves)", d->nOct);
    // Detector.cu:54-59 and :63-77
    float sigma_diff = sqrtf(fmax(c.sigma * c.sigma - 0.5f * 0.5f * 4, 0.01f));
    d->init_taps = gaussian_taps(sigma_diff);
    std::vector<float> sigmas(kGaussLayers);
    sigmas[0] = (float)c.sigma;
    double k = std::pow(2., 1. / c.numOctaveLayers);
    for (int i = 1; i < kGaussLayers; i++) {
        double sig_prev = std::pow(k, (double)(i - 1)) * c.sigma;
        double sig_total = sig_prev * k;
        sigmas[i] = (float)std::sqrt(sig_total * sig_total - sig_prev * sig_prev);
    }
    for (int i = 0; i < kGaussLayers; i++) d->taps[i] = gaussian_taps(sigmas[i]);
    if (!radius_supported(d->init_taps.n)) return fail(B200SIFT_EINVAL, "sigma %.3f: unsupported tap count %d", c.sigma, d->init_taps.n);
    for (int i = 1; i < kGaussLayers; i++)
        if (!radius_supported(d->taps[i].n)) return fail(B200SIFT_EINVAL, "sigma %.3f: unsupported tap count %d", c.sigma, d->taps[i].n);

    FrameParams& fp = d->fp;
    std::memset(&fp, 0, sizeof(fp));
    fp.nOct = d->nOct;
    fp.o_begin = 0;
    fp.o_end = d->nOct;
    fp.numFeatures = c.numFeatures;
    fp.extrema_thr = (float)std::floor(0.5 * c.contrastThreshould / c.numOctaveLayers * 255 * 1.f);  // Detector.cu:362
    fp.contrast_thr = (float)c.contrastThreshould;
    fp.edge_thr = (float)c.edgeThreshould;
    fp.sigma = (float)c.sigma;

    d->in_pitch = pitch_floats(c.col_width);
    d->u8_pitch = ((size_t)c.col_width + 511) / 512 * 512;
    for (int b = 0; b < 2; b++) {
        if (int rc = dev_alloc(d, &d->d_input[b], (size_t)d->in_pitch * c.row_width)) return rc;
        if (int rc = dev_alloc(d, &d->d_input_u8[b], d->u8_pitch * c.row_width)) return rc;
    }

    int rows_total = 0, scat_ctas = 0;
    for (int o = 0; o < d->nOct; o++) {
        OctParams& oc = fp.oct[o];
        oc.w = (int)roundf(c.col_width / powf(2, o));  // Detector.cu:83
        oc.h = (int)roundf(c.row_width / powf(2, o));
        if (oc.w < 1 || oc.h < 1) return fail(B200SIFT_EINVAL, "octave %d is empty", o);
        oc.pitch = pitch_floats(oc.w);
        oc.cap = std::min(oc.pitch * oc.h, c.numFeatures * 10);  // KeyPoint.cu:13, Detector.cu:109
        oc.words = 2 * ((oc.w + 63) / 64);  // two mask words (even / odd pixels) per 64-pixel strip
        oc.row0 = rows_total;
        rows_total += kScales * oc.h;
        oc.scat_cta0 = scat_ctas;
        scat_ctas += (kScales * oc.h + 7) / 8;
    }
    fp.rows_total = rows_total;
    if (int rc = dev_alloc(d, &d->d_rowcount_all, (size_t)rows_total)) return rc;
    d->rowcount_bytes = (size_t)rows_total * sizeof(int);
    int* d_counts = nullptr;
    if (int rc = dev_alloc(d, &d_counts, (size_t)4 * kMaxOctaves + 4)) return rc;
    fp.total = d_counts + 4 * kMaxOctaves;

    int task0 = 0;
    for (int o = 0; o < d->nOct; o++) {
        OctParams& oc = fp.oct[o];
        const size_t ls = (size_t)oc.pitch * oc.h;
        // +pitch floats of zeroed tail so that the reference's read one row past layer 5 of a
        // stale slot stays inside the allocation
        if (int rc = dev_alloc(d, &oc.gauss, ls * kGaussLayers + oc.pitch)) return rc;
        if (int rc = dev_alloc(d, &oc.mask, (size_t)kScales * oc.h * oc.words)) return rc;
        oc.rowcount = d->d_rowcount_all + oc.row0;
        if (int rc = dev_alloc(d, &oc.cand, (size_t)oc.cap)) return rc;
        if (int rc = dev_alloc(d, &oc.candF, (size_t)oc.cap)) return rc;
        if (int rc = dev_alloc(d, &oc.coll, (size_t)oc.cap)) return rc;
        if (int rc = dev_alloc(d, &oc.collF, (size_t)oc.cap)) return rc;
        if (int rc = dev_alloc(d, &oc.flag, (size_t)oc.cap)) return rc;
        if (int rc = dev_alloc(d, &oc.oricnt, (size_t)oc.cap)) return rc;
        oc.counts = d_counts + 4 * o;
        const int band = oc.h - 2 * kBorder;
        oc.det_cols = (oc.w > 2 * kBorder && band > 0) ? oc.words / 2 : 0;
        oc.det_rows = oc.det_cols ? (band + kDetRows - 1) / kDetRows : 0;
        oc.det_task0 = task0;
        task0 += oc.det_cols * oc.det_rows;
    }
    fp.det_tasks = task0;
    if (int rc = dev_alloc(d, &fp.out_kpts, (size_t)c.numFeatures)) return rc;
    if (int rc = dev_alloc(d, &fp.out_feats, (size_t)c.numFeatures)) return rc;
    if (int rc = dev_alloc(d, &fp.out_desc, (size_t)c.numFeatures * 128)) return rc;
    if (int rc = dev_alloc(d, &d->d_prev_desc, (size_t)c.numFeatures * 128)) return rc;
    d->match_scratch_bytes_ = match_scratch_bytes(c.numFeatures, c.numFeatures);
    {
        uint8_t* p = nullptr;
        if (int rc = dev_alloc(d, &p, d->match_scratch_bytes_)) return rc;
        d->match_scratch = p;
    }
    CU(cudaMallocHost(&d->h_total, sizeof(int)));
    *d->h_total = 0;
    CU(descriptor_setup());
    // everything but the descriptor launches (launch_descriptor) runs at the highest stream priority
    int prio_least = 0, prio_greatest = 0;
    CU(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    CU(cudaStreamCreateWithPriority(&d->stream, cudaStreamNonBlocking, prio_greatest));
    for (int i = 0; i < kSideStreams; i++) CU(cudaStreamCreateWithPriority(&d->side[i], cudaStreamNonBlocking, prio_greatest));
    for (int o = 0; o < kMaxOctaves; o++) {
        CU(cudaEventCreateWithFlags(&d->ev_fork[o], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&d->ev_join[o], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&d->ev_counts[o], cudaEventDisableTiming));
    }
    d->tail_first = d->nOct;
    for (int o = d->nOct - 1; o >= 0 && tail_octave(fp.oct[o].w, fp.oct[o].h); o--) d->tail_first = o;
    // graph branches over the multi-CTA octaves [0, tail_first): big octaves alone, the rest together
    d->n_side = 0;
    for (int o = 0; o < d->tail_first;) {
        int e = o + 1;
        const bool big = fp.oct[o].w * fp.oct[o].h >= kBranchPixels && d->n_side < kSideStreams - 1;
        if (!big) e = d->tail_first;
        d->br_begin[d->n_side] = o;
        d->br_end[d->n_side] = e;
        d->n_side++;
        o = e;
    }

    CU(cudaStreamCreateWithFlags(&d->copy_stream, cudaStreamNonBlocking));
    for (int b = 0; b < 2; b++) {
        CU(cudaEventCreateWithFlags(&d->ev_copied[b], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&d->ev_consumed[b], cudaEventDisableTiming));
    }
    // One graph for the whole frame (one instance per input buffer).
    for (int b = 0; b < 2; b++) {
        CU(cudaStreamBeginCapture(d->stream, cudaStreamCaptureModeThreadLocal));
        int rc = enqueue_frame(d, d->stream, 7, &d->launches, nullptr, b);
        cudaError_t e = cudaStreamEndCapture(d->stream, &d->graph[b]);
        if (rc) return rc;
        CU(e);
        CU(cudaGraphInstantiate(&d->exec[b], d->graph[b], 0));
    }
    CU(cudaStreamSynchronize(d->stream));
    return 0;
}

}  // namespace

// ------------------------------------------------------------------------------------ C ABI
extern "C" {

void b200sift_default_config(b200sift_config* cfg) {
    if (!cfg) return;
    cfg->col_width = 0;
    cfg->row_width = 0;
    cfg->numFeatures = 5000;
    cfg->numOctaveLayers = 3;
    cfg->contrastThreshould = 0.04;
    cfg->edgeThreshould = 10;
    cfg->sigma = 1.6;
    cfg->upscale = 0;
}

const char* b200sift_last_error(void) { return g_err.c_str(); }
const char* b200sift_version(void) { return "b200sift 0.1 (sm_100a)"; }

int b200sift_create(const b200sift_config* cfg, b200sift_detector** out) {
    if (!cfg || !out) return fail(B200SIFT_EINVAL, "null argument");
    *out = nullptr;
    if (cfg->col_width <= 0 || cfg->row_width <= 0)
        return fail(B200SIFT_EINVAL, "Image width or height not set.");  // Detector.cu:22-25
    if (cfg->upscale) return fail(B200SIFT_EINVAL, "upscale=true is broken in the reference (CudaSiftConfig.hh:12) and unsupported");
    if (cfg->numOctaveLayers != 3) return fail(B200SIFT_EINVAL, "numOctaveLayers must be 3");
    if (cfg->numFeatures <= 0) return fail(B200SIFT_EINVAL, "numFeatures must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(B200SIFT_ECUDA, "no CUDA device: this engine has no CPU fallback");
    }
    b200sift_detector* d = new b200sift_detector;
    d->cfg = *cfg;
    int rc = build(d);
    if (rc) {
        std::string keep = g_err;
        b200sift_destroy(d);
        g_err = keep;
        return rc;
    }
    *out = d;
    return 0;
}

int b200sift_destroy(b200sift_detector* d) {
    if (!d) return 0;
    if (d->stream) cudaStreamSynchronize(d->stream);
    for (int b = 0; b < 2; b++) {
        if (d->exec[b]) cudaGraphExecDestroy(d->exec[b]);
        if (d->graph[b]) cudaGraphDestroy(d->graph[b]);
        if (d->ev_copied[b]) cudaEventDestroy(d->ev_copied[b]);
        if (d->ev_consumed[b]) cudaEventDestroy(d->ev_consumed[b]);
    }
    if (d->copy_stream) cudaStreamDestroy(d->copy_stream);
    for (void* p : d->allocs) cudaFree(p);
    if (d->h_total) cudaFreeHost(d->h_total);
    if (d->stream) cudaStreamDestroy(d->stream);
    for (int i = 0; i < kSideStreams; i++)
        if (d->side[i]) cudaStreamDestroy(d->side[i]);
    for (int o = 0; o < kMaxOctaves; o++) {
        if (d->ev_fork[o]) cudaEventDestroy(d->ev_fork[o]);
        if (d->ev_counts[o]) cudaEventDestroy(d->ev_counts[o]);
        if (d->ev_join[o]) cudaEventDestroy(d->ev_join[o]);
    }
    delete d;
    return 0;
}

int b200sift_num_octaves(const b200sift_detector* d) { return d ? d->nOct : B200SIFT_EINVAL; }

int b200sift_octave_info(const b200sift_detector* d, int o, int* dims) {
    if (!d || !dims || o < 0 || o >= d->nOct) return fail(B200SIFT_EINVAL, "bad octave");
    const OctParams& oc = d->fp.oct[o];
    dims[0] = oc.w; dims[1] = oc.h; dims[2] = oc.pitch; dims[3] = oc.cap;
    return 0;
}

// Frame i uses input buffer i & 1. Host frames are copied on the copy stream, which only waits until the
// frame that last read this buffer (i - 2) has consumed it, so the copy overlaps the graph of frame i - 1.
static int begin_frame(b200sift_detector* d, int* buf, bool on_copy_stream) {
    *buf = (int)(d->frame_no & 1);
    if (on_copy_stream) CU(cudaStreamWaitEvent(d->copy_stream, d->ev_consumed[*buf], 0));
    return 0;
}
static int launch_frame(b200sift_detector* d, int buf, bool from_copy_stream) {
    if (from_copy_stream) {
        CU(cudaEventRecord(d->ev_copied[buf], d->copy_stream));
        CU(cudaStreamWaitEvent(d->stream, d->ev_copied[buf], 0));
    }
    CU(cudaGraphLaunch(d->exec[buf], d->stream));
    CU(cudaEventRecord(d->ev_consumed[buf], d->stream));
    d->frame_no++;
    d->have_frame = true;
    return 0;
}

int b200sift_detect_host_async(b200sift_detector* d, const float* img, int rows, int cols) {
    if (!d || !img) return fail(B200SIFT_EINVAL, "null argument");
    if (rows != d->cfg.row_width || cols != d->cfg.col_width)
        return fail(B200SIFT_EINVAL, "image is %dx%d, detector was configured for %dx%d", cols, rows, d->cfg.col_width, d->cfg.row_width);
    int buf = 0;
    if (int rc = begin_frame(d, &buf, true)) return rc;
    CU(cudaMemcpy2DAsync(d->d_input[buf], (size_t)d->in_pitch * 4, img, (size_t)cols * 4, (size_t)cols * 4, rows,
                         cudaMemcpyHostToDevice, d->copy_stream));
    return launch_frame(d, buf, true);
}

int b200sift_sync(b200sift_detector* d, int* total) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    CU(cudaStreamSynchronize(d->stream));
    if (total) *total = d->have_frame ? *d->h_total : 0;
    return 0;
}

int b200sift_detect_host(b200sift_detector* d, const float* img, int rows, int cols, int* total) {
    int rc = b200sift_detect_host_async(d, img, rows, cols);
    if (rc) return rc;
    return b200sift_sync(d, total);
}

int b200sift_detect_device_async(b200sift_detector* d, const float* d_img, size_t pitch_bytes) {
    if (!d || !d_img) return fail(B200SIFT_EINVAL, "null argument");
    const int cols = d->cfg.col_width, rows = d->cfg.row_width;
    if (pitch_bytes < (size_t)cols * 4) return fail(B200SIFT_EINVAL, "pitch smaller than a row");
    int buf = 0;
    if (int rc = begin_frame(d, &buf, false)) return rc;
    CU(cudaMemcpy2DAsync(d->d_input[buf], (size_t)d->in_pitch * 4, d_img, pitch_bytes, (size_t)cols * 4, rows,
                         cudaMemcpyDeviceToDevice, d->stream));
    return launch_frame(d, buf, false);
}

int b200sift_detect_host_u8_async(b200sift_detector* d, const uint8_t* img, int rows, int cols) {
    if (!d || !img) return fail(B200SIFT_EINVAL, "null argument");
    if (rows != d->cfg.row_width || cols != d->cfg.col_width)
        return fail(B200SIFT_EINVAL, "image is %dx%d, detector was configured for %dx%d", cols, rows, d->cfg.col_width, d->cfg.row_width);
    int buf = 0;
    if (int rc = begin_frame(d, &buf, true)) return rc;
    CU(cudaMemcpy2DAsync(d->d_input_u8[buf], d->u8_pitch, img, (size_t)cols, (size_t)cols, rows, cudaMemcpyHostToDevice, d->copy_stream));
    launch_u8_to_float(d->d_input_u8[buf], d->u8_pitch, d->d_input[buf], d->in_pitch, cols, rows, d->copy_stream);
    return launch_frame(d, buf, true);
}

int b200sift_match_consecutive_async(b200sift_detector* d, int* d_out_idx) {
    if (!d || !d_out_idx) return fail(B200SIFT_EINVAL, "null argument");
    if (match_launch_counts(d->d_prev_desc, d->fp.out_desc, d->cfg.numFeatures, d->fp.total, d_out_idx, d->match_scratch,
                            d->match_scratch_bytes_, d->stream))
        return fail(B200SIFT_EINVAL, "matcher scratch too small");
    CU(cudaGetLastError());
    return 0;
}

int b200sift_copy_results_async(b200sift_detector* d, float* kpts3, float* feats4, uint16_t* desc, int* counts2) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    const size_t n = (size_t)d->cfg.numFeatures;
    if (kpts3) CU(cudaMemcpyAsync(kpts3, d->fp.out_kpts, n * sizeof(float3), cudaMemcpyDeviceToHost, d->stream));
    if (feats4) CU(cudaMemcpyAsync(feats4, d->fp.out_feats, n * sizeof(float4), cudaMemcpyDeviceToHost, d->stream));
    if (desc) CU(cudaMemcpyAsync(desc, d->fp.out_desc, n * 128 * sizeof(__half), cudaMemcpyDeviceToHost, d->stream));
    if (counts2) CU(cudaMemcpyAsync(counts2, d->fp.total, 2 * sizeof(int), cudaMemcpyDeviceToHost, d->stream));
    return 0;
}

void* b200sift_stream(b200sift_detector* d) { return d ? (void*)d->stream : nullptr; }

int b200sift_results(b200sift_detector* d, const float** kpts, const float** feats, const uint16_t** desc,
                     const uint16_t** prev_desc, int* total) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    if (kpts) *kpts = reinterpret_cast<const float*>(d->fp.out_kpts);
    if (feats) *feats = reinterpret_cast<const float*>(d->fp.out_feats);
    if (desc) *desc = reinterpret_cast<const uint16_t*>(d->fp.out_desc);
    if (prev_desc) *prev_desc = reinterpret_cast<const uint16_t*>(d->d_prev_desc);
    if (total) return b200sift_sync(d, total);
    return 0;
}

int b200sift_copy_to_host(b200sift_detector* d, float* kpts3, float* feats4, uint16_t* desc, int with_desc) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    int total = 0;
    int rc = b200sift_sync(d, &total);
    if (rc) return rc;
    if (kpts3) CU(cudaMemcpyAsync(kpts3, d->fp.out_kpts, (size_t)total * sizeof(float3), cudaMemcpyDeviceToHost, d->stream));
    if (feats4) CU(cudaMemcpyAsync(feats4, d->fp.out_feats, (size_t)total * sizeof(float4), cudaMemcpyDeviceToHost, d->stream));
    if (with_desc && desc)
        CU(cudaMemcpyAsync(desc, d->fp.out_desc, (size_t)total * 128 * sizeof(__half), cudaMemcpyDeviceToHost, d->stream));
    CU(cudaStreamSynchronize(d->stream));
    return 0;
}

int b200sift_debug_get_stack(b200sift_detector* d, int o, int which, float* out) {
    if (!d || !out || o < 0 || o >= d->nOct || which != 0) return fail(B200SIFT_EINVAL, "bad argument");
    const OctParams& oc = d->fp.oct[o];
    CU(cudaStreamSynchronize(d->stream));
    CU(cudaMemcpy2D(out, (size_t)oc.w * 4, oc.gauss, (size_t)oc.pitch * 4, (size_t)oc.w * 4, (size_t)oc.h * kGaussLayers,
                    cudaMemcpyDeviceToHost));
    return 0;
}

int b200sift_debug_get_kpt_buffer(b200sift_detector* d, int o, int which, void* out, int max_elems) {
    if (!d || !out || o < 0 || o >= d->nOct) return fail(B200SIFT_EINVAL, "bad argument");
    const OctParams& oc = d->fp.oct[o];
    const int n = std::min(max_elems, oc.cap);
    CU(cudaStreamSynchronize(d->stream));
    switch (which) {
        case 0: CU(cudaMemcpy(out, oc.cand, (size_t)n * sizeof(float3), cudaMemcpyDeviceToHost)); break;
        case 1: CU(cudaMemcpy(out, oc.candF, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost)); break;
        case 2: CU(cudaMemcpy(out, oc.coll, (size_t)n * sizeof(float3), cudaMemcpyDeviceToHost)); break;
        case 3: CU(cudaMemcpy(out, oc.collF, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost)); break;
        default: return fail(B200SIFT_EINVAL, "bad buffer id");
    }
    return oc.cap;
}

int b200sift_debug_get_counts(b200sift_detector* d, int o, int* counts) {
    if (!d || !counts || o < 0 || o >= d->nOct) return fail(B200SIFT_EINVAL, "bad argument");
    CU(cudaStreamSynchronize(d->stream));
    int c[4];
    CU(cudaMemcpy(c, d->fp.oct[o].counts, sizeof(c), cudaMemcpyDeviceToHost));
    counts[0] = c[0]; counts[1] = c[1]; counts[2] = c[2];
    return 0;
}

int b200sift_debug_get_taps(b200sift_detector* d, int which, float* out, int max_elems) {
    if (!d || which < -1 || which >= kGaussLayers) return fail(B200SIFT_EINVAL, "bad argument");
    const Taps& t = which < 0 ? d->init_taps : d->taps[which];
    if (out && t.n <= max_elems) std::memcpy(out, t.v, t.n * sizeof(float));
    return t.n;
}

int b200sift_debug_run_stages(b200sift_detector* d, const float* img, int rows, int cols, int last_stage) {
    if (!d || !img) return fail(B200SIFT_EINVAL, "null argument");
    if (rows != d->cfg.row_width || cols != d->cfg.col_width) return fail(B200SIFT_EINVAL, "image size mismatch");
    CU(cudaStreamSynchronize(d->stream));
    const int buf = (int)(d->frame_no & 1);
    CU(cudaMemcpy2DAsync(d->d_input[buf], (size_t)d->in_pitch * 4, img, (size_t)cols * 4, (size_t)cols * 4, rows,
                         cudaMemcpyHostToDevice, d->stream));
    int rc = enqueue_frame(d, d->stream, last_stage, nullptr, nullptr, buf);
    if (rc) return rc;
    if (last_stage >= 6) d->have_frame = true;
    CU(cudaStreamSynchronize(d->stream));
    return 0;
}

int b200sift_launches_per_frame(const b200sift_detector* d) { return d ? d->launches : B200SIFT_EINVAL; }

int b200sift_time_frames(b200sift_detector* d, int warmup, int iters, float* ms_per_frame) {
    if (!d || !ms_per_frame || iters <= 0) return fail(B200SIFT_EINVAL, "bad argument");
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int tb = (int)((d->frame_no + 1) & 1);  // buffer of the last frame
    for (int i = 0; i < warmup; i++) CU(cudaGraphLaunch(d->exec[tb], d->stream));
    CU(cudaStreamSynchronize(d->stream));
    CU(cudaEventRecord(e0, d->stream));
    for (int i = 0; i < iters; i++) CU(cudaGraphLaunch(d->exec[tb], d->stream));
    CU(cudaEventRecord(e1, d->stream));
    CU(cudaEventSynchronize(e1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    d->have_frame = true;
    *ms_per_frame = ms / (float)iters;
    return 0;
}

int b200sift_profile_frame(b200sift_detector* d, int warmup, int iters, float* ms_per_class, int* launches_per_class) {
    if (!d || !ms_per_class || iters <= 0) return fail(B200SIFT_EINVAL, "bad argument");
    Profiler prof;
    prof.s = d->stream;
    std::vector<double> acc(B200SIFT_K_COUNT, 0.0);
    std::vector<int> cnt(B200SIFT_K_COUNT, 0);
    for (int it = 0; it < warmup + iters; it++) {
        prof.used = 0;
        int rc = enqueue_frame(d, d->stream, 7, nullptr, &prof, (int)((d->frame_no + 1) & 1));
        if (rc) return rc;
        CU(cudaStreamSynchronize(d->stream));
        if (it < warmup) continue;
        for (size_t i = 1; i < prof.used; i++) {
            float ms = 0.f;
            CU(cudaEventElapsedTime(&ms, prof.ev[i - 1], prof.ev[i]));
            acc[prof.cls[i]] += ms;
            if (it == warmup) cnt[prof.cls[i]]++;
        }
    }
    d->have_frame = true;
    for (int c = 0; c < B200SIFT_K_COUNT; c++) {
        ms_per_class[c] = (float)(acc[c] / iters);
        if (launches_per_class) launches_per_class[c] = cnt[c];
    }
    return 0;
}

// In-graph cumulative stage times: the frame graph truncated after the pyramid (1), extrema (3), refinement
// (4), orientation (5) and complete (7), each captured and replayed `iters` times on the last frame's input.
int b200sift_time_stages(b200sift_detector* d, int iters, float* ms5) {
    if (!d || !ms5 || iters <= 0) return fail(B200SIFT_EINVAL, "bad argument");
    const int stages[5] = {1, 3, 4, 5, 7};
    const int buf = (int)((d->frame_no + 1) & 1);
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    for (int i = 0; i < 5; i++) {
        cudaGraph_t g = nullptr;
        cudaGraphExec_t x = nullptr;
        CU(cudaStreamSynchronize(d->stream));
        CU(cudaStreamBeginCapture(d->stream, cudaStreamCaptureModeThreadLocal));
        int rc = enqueue_frame(d, d->stream, stages[i], nullptr, nullptr, buf);
        cudaError_t e = cudaStreamEndCapture(d->stream, &g);
        if (rc) return rc;
        CU(e);
        CU(cudaGraphInstantiate(&x, g, 0));
        for (int k = 0; k < 5; k++) CU(cudaGraphLaunch(x, d->stream));
        CU(cudaEventRecord(e0, d->stream));
        for (int k = 0; k < iters; k++) CU(cudaGraphLaunch(x, d->stream));
        CU(cudaEventRecord(e1, d->stream));
        CU(cudaEventSynchronize(e1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        ms5[i] = ms / (float)iters;
        cudaGraphExecDestroy(x);
        cudaGraphDestroy(g);
    }
    // leave complete results behind
    CU(cudaGraphLaunch(d->exec[buf], d->stream));
    CU(cudaStreamSynchronize(d->stream));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return 0;
}

// ---- matcher
namespace {
std::mutex g_scratch_mu;
void* g_scratch[64] = {nullptr};
size_t g_scratch_bytes[64] = {0};
}  // namespace

int b200sift_match_device(const uint16_t* d_query, int nq, const uint16_t* d_train, int nt, int* d_out_idx, void* stream) {
    if (nq < 0 || nt < 0 || (nq > 0 && (!d_query || !d_out_idx)) || (nt > 0 && !d_train))
        return fail(B200SIFT_EINVAL, "bad argument");
    if (nq == 0) return 0;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(B200SIFT_EINVAL, "device index out of range");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const size_t need = match_scratch_bytes(nq, nt);
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_scratch_mu);
        if (g_scratch_bytes[dev] < need) {
            if (g_scratch[dev]) {
                CU(cudaDeviceSynchronize());
                CU(cudaFree(g_scratch[dev]));
                g_scratch[dev] = nullptr;
                g_scratch_bytes[dev] = 0;
            }
            CU(cudaMalloc(&g_scratch[dev], need));
            g_scratch_bytes[dev] = need;
        }
        scratch = g_scratch[dev];
    }
    if (match_launch(reinterpret_cast<const __half*>(d_query), nq, reinterpret_cast<const __half*>(d_train), nt, d_out_idx,
                     scratch, need, s))
        return fail(B200SIFT_EINVAL, "matcher scratch too small");
    CU(cudaGetLastError());
    return 0;
}

int b200sift_match_host(const uint16_t* query, int nq, const uint16_t* train, int nt, int* out_idx) {
    if (nq < 0 || nt < 0 || (nq > 0 && (!query || !out_idx)) || (nt > 0 && !train)) return fail(B200SIFT_EINVAL, "bad argument");
    if (nq == 0) return 0;
    uint16_t *dq = nullptr, *dt = nullptr;
    int* dout = nullptr;
    CU(cudaMalloc(&dq, (size_t)nq * 256));
    CU(cudaMalloc(&dt, std::max<size_t>((size_t)nt * 256, 16)));
    CU(cudaMalloc(&dout, (size_t)nq * sizeof(int)));
    CU(cudaMemcpy(dq, query, (size_t)nq * 256, cudaMemcpyHostToDevice));
    if (nt) CU(cudaMemcpy(dt, train, (size_t)nt * 256, cudaMemcpyHostToDevice));
    int rc = b200sift_match_device(dq, nq, dt, nt, dout, nullptr);
    if (!rc) {
        cudaError_t e = cudaMemcpy(out_idx, dout, (size_t)nq * sizeof(int), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(B200SIFT_ECUDA, "%s", cudaGetErrorString(e));
    }
    cudaFree(dq);
    cudaFree(dt);
    cudaFree(dout);
    return rc;
}

int b200sift_match_launches(int nq, int nt) { return match_num_launches(nq, nt); }

int b200sift_match_set_path(int path) {
    if (path != 0 && path != 1) return fail(B200SIFT_EINVAL, "path must be 0 (tensor core) or 1 (CUDA core)");
    match_set_path(path);
    return 0;
}

int b200sift_debug_match_dist(const uint16_t* d_query, int nq, const uint16_t* d_train, int nt, float* d_dist, int* d_out_idx) {
    if (nq <= 0 || nt <= 0 || !d_query || !d_train || !d_dist || !d_out_idx) return fail(B200SIFT_EINVAL, "bad argument");
    void* scratch = nullptr;
    const size_t need = match_scratch_bytes(nq, nt);
    CU(cudaMalloc(&scratch, need));
    int rc = match_launch(reinterpret_cast<const __half*>(d_query), nq, reinterpret_cast<const __half*>(d_train), nt, d_out_idx,
                          scratch, need, nullptr, d_dist);
    cudaError_t e = cudaDeviceSynchronize();
    cudaFree(scratch);
    if (rc) return fail(B200SIFT_EINVAL, "matcher scratch too small");
    CU(e);
    return 0;
}

}  // extern "C"
