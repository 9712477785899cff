// Host side of the B200 SIFT engine: allocation, per-frame CUDA graph, C ABI (include/b200sift.h).
//
// Replaces the orchestration of interface/Detector.cu (5 CUDA graphs, 6 host synchronisations and
// a pinned-memory count round trip per frame) with ONE graph per frame and no host involvement
// between its first and last kernel: every count stays on the device.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
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

}  // namespace

namespace b200sift {
// last-error setter for the other translation units behind the C ABI (shard.cu)
int set_error(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
}  // namespace b200sift

namespace {

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

bool taps_supported(const Taps& t) { return (t.n & 1) && t.n >= 3 && t.n <= kMaxTaps; }

inline int pitch_floats(int w) { return ((w * 4 + 511) / 512) * 512 / 4; }

}  // namespace

constexpr int kSideStreams = 6;   // side branches of the frame graph
constexpr int kBranchPixels = 65536;  // an octave with at least this many pixels gets a branch of its own

// Frames in flight: a create-time option (b200sift_options::frames_in_flight, 3..32), ~85 MB of device memory per slot
// at 1920x1200. Every slot has its own stream, so frame n + nslots queues behind frame n: with few slots the pyramid of a
// new frame cannot start before the descriptors of the frame that held its slot are done. Measured on one B200 (same box,
// frames/s with device-resident input, ~3800 keypoints per frame): 6 -> 3772, 8 -> 3948, 12 -> 3975, 16 -> 4147, 32 -> 4209
// at 1920x1200; 752x480: 6 -> 8920, 16 -> 10979 (profiles/r2_pyramid_ab.md section 4).
constexpr int kDefaultSlots = 16, kMinSlots = 3, kMaxSlots = 32;

// Everything one frame touches. Frame n runs in slot n % nslots on the slot's own streams, so the frame graph
// of frame n+1 (pyramid, extrema: latency-bound chains of small launches) overlaps the descriptor kernels of
// frame n, and the consecutive-frame matcher can read the outputs of frames n-1 and n in place while frame
// n+1 is already running: with three slots nothing is copied between frames.
struct Slot {
    FrameParams fp;
    float* d_input = nullptr;
    float* d_up = nullptr;           // the doubled frame (working upscale only)
    uint8_t* d_input_u8 = nullptr;
    int* d_rowcount_all = nullptr;
    int* h_total = nullptr;          // pinned
    void* match_scratch = nullptr;   // survivor records of the consecutive-frame matcher
    int* d_match_idx = nullptr;      // match indices of b200sift_detect_batch_host (numFeatures ints)
    cudaStream_t stream = nullptr;   // main stream of the slot's frame graph
    cudaStream_t side[kSideStreams] = {nullptr};
    cudaStream_t stream_a = nullptr;  // capture-time streams of part A (their priority ends up in its kernel nodes)
    cudaStream_t side_a[kSideStreams] = {nullptr};
    cudaEvent_t ev_fork[kMaxOctaves] = {nullptr}, ev_join[kMaxOctaves] = {nullptr}, ev_counts[kMaxOctaves] = {nullptr};
    cudaEvent_t ev_copied = nullptr, ev_consumed = nullptr;  // input buffer hand-over with the copy stream
    cudaEvent_t ev_done = nullptr;      // the slot's frame is complete
    cudaEvent_t ev_matched = nullptr;   // the last matcher that read this slot as "previous frame" is complete
    bool matched_pending = false;
    cudaEvent_t ev_chain = nullptr;     // part B (the shared candidate lists) of the slot's frame is complete
    cudaGraph_t graph[3] = {nullptr, nullptr, nullptr};       // parts A, B, C of the frame (enqueue_frame)
    cudaGraphExec_t exec[3] = {nullptr, nullptr, nullptr};
};

struct b200sift_detector {
    b200sift_config cfg;
    int device = 0;
    int nOct = 0;
    int nScales = 3, nLayers = 6;     // numOctaveLayers, numOctaveLayers + 3
    Taps init_taps;
    Taps taps[kMaxLayers];
    Slot slot[kMaxSlots];
    int nslots = kDefaultSlots;
    b200sift_options opt;
    int in_pitch = 0;
    bool upscaled = false;            // cfg.upscale with b200sift_options::working_upscale
    int base_w = 0, base_h = 0, up_pitch = 0;   // dimensions of octave 0 (2x the frame when upscaled)
    size_t u8_pitch = 0;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_user = nullptr;    // orders a caller's stream before the engine's use of a caller-owned device buffer
    unsigned long long frame_no = 0;  // frames enqueued so far; frame n lives in slot n % kSlots
    size_t rowcount_bytes = 0;
    size_t match_scratch_bytes_ = 0;
    std::vector<void*> allocs;
    int tail_first = 0;  // first octave handled by the single-CTA tail kernel
    // Branches of the frame graph: octave ranges [br_begin, br_end). Branches 0 .. n_side-1 cover the
    // multi-CTA octaves and run on side streams; the tail octaves (if any) run on the main stream.
    int n_side = 0;
    int br_begin[kSideStreams] = {0}, br_end[kSideStreams] = {0};
    int launches = 0;
    bool have_frame = false;
    Slot& cur() { return slot[(frame_no + nslots - 1) % nslots]; }    // slot of the latest frame
    Slot& prev() { return slot[(frame_no + nslots - 2) % nslots]; }   // slot of the frame before it
    const Slot& cur() const { return slot[(frame_no + nslots - 1) % nslots]; }
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
// 1 pyramid, 3 extrema, 4 refinement, 5 orientation, 6/7 descriptor+gather. It is enqueued (and captured) in
// three parts, because the candidate lists (cand/candF/coll/collF) are shared by the frame slots -- the
// reference leaves stale entries in them from frame to frame (quirk Q5) and the results depend on it:
//   A  pyramid + extrema bitmask        per-slot buffers only: overlaps anything of the previous frames
//   B  scatter .. orientation           touches the shared lists: runs strictly in frame order
//   C  descriptors                      per-slot buffers only (B leaves the final list in fin_kp/fin_ft)
enum : int { kPartA = 1, kPartB = 2, kPartC = 4, kPartAll = 7 };

int enqueue_frame(b200sift_detector* d, Slot& sl, int last_stage, int* launches, Profiler* prof = nullptr, int parts = kPartAll) {
    FrameParams& fp = sl.fp;
    cudaStream_t s = sl.stream;
    int n = 0;
    const bool fork = (prof == nullptr) && sl.side[0] != nullptr;
    bool used[kSideStreams] = {false};
    auto join_sides = [&]() -> int {
        for (int b = 0; b < d->n_side; b++)
            if (used[b]) {
                CU(cudaEventRecord(sl.ev_join[b], sl.side[b]));
                CU(cudaStreamWaitEvent(s, sl.ev_join[b], 0));
                used[b] = false;
            }
        return 0;
    };

    if (parts & kPartA) {
        MARK(B200SIFT_K_COPY);
        CU(cudaMemsetAsync(sl.d_rowcount_all, 0, d->rowcount_bytes, s));
        MARK(B200SIFT_K_COPY);
        // Gaussian pyramid. Octave o+1 only needs layer 3 of octave o, so layers 4 and 5 of every big octave
        // go to a side stream (a parallel branch of the graph) while the main chain moves on to the next
        // octave; octaves of <= kTailPixels pixels are walked by one CTA (launch_pyramid_tail). The graph
        // branches per octave range (build(): one branch per big octave, one for the remaining multi-CTA
        // octaves, the tail octaves on the main stream); each branch runs its extrema detection as soon as
        // the six layers of its octaves exist.
        // Launches per octave of the default configuration: octave 0 = one single-stage launch per layer (initial blur,
        // L1..Ln on the main chain, L(n+1), L(n+2) on the branch); octaves >= 1 = ONE fused launch on the main chain
        // (2x resample of the previous octave's layer n + L1..Ln) and ONE on the branch (L(n+1), L(n+2)). Radii without a
        // fused instantiation fall back to resample + one launch per layer.
        const int nL = d->nLayers, nS = d->nScales;
        int br = 0;  // branch of the current multi-CTA octave
        for (int o = 0; o < d->tail_first; o++) {
            OctParams& oc = fp.oct[o];
            const size_t ls = (size_t)oc.pitch * oc.h;
            const int cls = o == 0 ? B200SIFT_K_BLUR : B200SIFT_K_BLUR_SMALL;
            BlurStage st[kMaxLayers];
            for (int l = 1; l < nL; l++) st[l] = BlurStage{oc.gauss + ls * l, &d->taps[l]};
            bool fused_main = false;
            if (o == 0) {
                const float* base = sl.d_input;
                int base_pitch = d->in_pitch;
                if (d->upscaled) {  // working upscale (opt-in): the frame doubled with the reference's bilinear resize (Detector.cu:236-244)
                    launch_downsample(sl.d_input, d->in_pitch, d->cfg.col_width, d->cfg.row_width, sl.d_up, d->up_pitch, oc.w, oc.h, s);
                    MARK(B200SIFT_K_DOWNSAMPLE);
                    n++;
                    base = sl.d_up;
                    base_pitch = d->up_pitch;
                }
                launch_blur(base, base_pitch, oc.gauss, oc.pitch, oc.w, oc.h, d->init_taps, s);
                MARK(B200SIFT_K_BLUR);
                n++;
            } else {
                const OctParams& pv = fp.oct[o - 1];
                const float* src = pv.gauss + (size_t)pv.pitch * pv.h * nS;
                fused_main = nS <= 3 && blur_chain_supported(true, st + 1, nS);
                if (fused_main) {
                    launch_blur_chain(src, pv.pitch, pv.w, pv.h, true, oc.gauss, st + 1, nS, oc.pitch, oc.w, oc.h, s);
                    MARK(cls);
                } else {
                    launch_downsample(src, pv.pitch, pv.w, pv.h, oc.gauss, oc.pitch, oc.w, oc.h, s);
                    MARK(B200SIFT_K_DOWNSAMPLE);
                }
                n++;
            }
            if (!fused_main)
                for (int l = 1; l <= nS; l++) {
                    launch_blur(oc.gauss + ls * (l - 1), oc.pitch, st[l].out, oc.pitch, oc.w, oc.h, d->taps[l], s);
                    MARK(cls);
                    n++;
                }
            while (o >= d->br_end[br]) br++;
            cudaStream_t bs = fork ? sl.side[br] : s;  // the branch's stream
            if (fork) {  // layers n+1, n+2 off the critical path
                CU(cudaEventRecord(sl.ev_fork[o], s));
                CU(cudaStreamWaitEvent(bs, sl.ev_fork[o], 0));
                used[br] = true;
            }
            if (o > 0 && blur_chain_supported(false, st + nS + 1, 2)) {
                launch_blur_chain(oc.gauss + ls * nS, oc.pitch, oc.w, oc.h, false, nullptr, st + nS + 1, 2, oc.pitch, oc.w, oc.h, bs);
                MARK(cls);
                n++;
            } else {
                for (int l = nS + 1; l < nL; l++) {
                    launch_blur(oc.gauss + ls * (l - 1), oc.pitch, st[l].out, oc.pitch, oc.w, oc.h, d->taps[l], bs);
                    MARK(cls);
                    n++;
                }
            }
            if (o == d->br_end[br] - 1 && last_stage >= 3) {  // the branch's octaves are complete
                launch_detect(fp, d->br_begin[br], d->br_end[br], bs); n++;
                MARK(B200SIFT_K_DETECT);
            }
        }
        if (d->tail_first < d->nOct) {
            // tail octaves: one CTA walks resample + L0..Ln of all of them (a dependency chain), then one CTA per octave
            // adds L(n+1), L(n+2)
            const float* base = d->upscaled ? sl.d_up : sl.d_input;
            const int base_pitch = d->upscaled ? d->up_pitch : d->in_pitch;
            if (d->upscaled && d->tail_first == 0) {
                launch_downsample(sl.d_input, d->in_pitch, d->cfg.col_width, d->cfg.row_width, sl.d_up, d->up_pitch, fp.oct[0].w, fp.oct[0].h, s);
                MARK(B200SIFT_K_DOWNSAMPLE);
                n++;
            }
            launch_pyramid_tail(fp, d->taps, d->init_taps, base, base_pitch, d->tail_first, 0, nS + 1, s);
            MARK(B200SIFT_K_BLUR_SMALL);
            launch_pyramid_tail(fp, d->taps, d->init_taps, base, base_pitch, d->tail_first, nS + 1, nL, s);
            MARK(B200SIFT_K_BLUR_SMALL);
            n += 2;
        }
        if (d->tail_first < d->nOct && last_stage >= 3) {  // tail octaves on the main stream
            launch_detect(fp, d->tail_first, d->nOct, s); n++;
            MARK(B200SIFT_K_DETECT);
        }
        if (int rc = join_sides()) return rc;
    }

    if ((parts & kPartB) && last_stage >= 3) {
        auto chain = [&](int ob, int oe, cudaStream_t cs) {
            launch_scatter(fp, ob, oe, cs); n++;
            MARK(B200SIFT_K_COMPACT);
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
                if (fp.opt_dedup) {
                    launch_dedup(fp, ob, oe, cs); n += 3;
                    MARK(B200SIFT_K_COMPACT);
                }
            }
        };
        if (fork && d->n_side > 0) CU(cudaEventRecord(sl.ev_fork[0], s));
        for (int b = 0; b < d->n_side; b++) {
            cudaStream_t bs = fork ? sl.side[b] : s;
            if (fork) {
                CU(cudaStreamWaitEvent(bs, sl.ev_fork[0], 0));
                used[b] = true;
            }
            chain(d->br_begin[b], d->br_end[b], bs);
        }
        if (d->tail_first < d->nOct) chain(d->tail_first, d->nOct, s);
        if (int rc = join_sides()) return rc;
    }

    if ((parts & kPartC) && last_stage >= 6) {
        CU(launch_descriptor(fp, 0, d->nOct, s)); n++;
        MARK(B200SIFT_K_DESCRIBE);
        CU(cudaMemcpyAsync(sl.h_total, fp.total, sizeof(int), cudaMemcpyDeviceToHost, s));
        MARK(B200SIFT_K_COPY);
    }
    CU(cudaGetLastError());
    if (pyramid_failed()) return fail(B200SIFT_ECUDA, "cuTensorMapEncodeTiled rejected a Gaussian layer (driver too old for TMA?)");
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
    d->nScales = c.numOctaveLayers;
    d->nLayers = c.numOctaveLayers + 3;
    const int nL = d->nLayers, nS = d->nScales;
    std::vector<float> sigmas(nL);
    sigmas[0] = (float)c.sigma;
    double k = std::pow(2., 1. / c.numOctaveLayers);
    for (int i = 1; i < nL; i++) {
        double sig_prev = std::pow(k, (double)(i - 1)) * c.sigma;
        double sig_total = sig_prev * k;
        sigmas[i] = (float)std::sqrt(sig_total * sig_total - sig_prev * sig_prev);
    }
    for (int i = 0; i < nL; i++) d->taps[i] = gaussian_taps(sigmas[i]);
    if (!taps_supported(d->init_taps))
        return fail(B200SIFT_EINVAL, "sigma %.3f: the initial blur needs %d taps (supported: 3..%d)", c.sigma, d->init_taps.n, kMaxTaps);
    for (int i = 1; i < nL; i++)
        if (!taps_supported(d->taps[i]))
            return fail(B200SIFT_EINVAL, "sigma %.3f, numOctaveLayers %d: layer %d needs %d taps (supported: 3..%d)", c.sigma,
                        c.numOctaveLayers, i, d->taps[i].n, kMaxTaps);

    d->in_pitch = pitch_floats(c.col_width);
    d->upscaled = c.upscale != 0;   // b200sift_create_ex only lets it through with working_upscale
    d->base_w = d->upscaled ? 2 * c.col_width : c.col_width;   // Detector.cu:44-50
    d->base_h = d->upscaled ? 2 * c.row_width : c.row_width;
    d->up_pitch = pitch_floats(d->base_w);
    d->u8_pitch = ((size_t)c.col_width + 511) / 512 * 512;
    d->match_scratch_bytes_ = match_scratch_bytes(c.numFeatures, c.numFeatures);
    CU(descriptor_setup());
    CU(pyramid_setup());
    // everything but the descriptor launches (launch_descriptor) runs at the highest stream priority
    int prio_least = 0, prio_greatest = 0;
    CU(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    int prio_a = prio_greatest;
    if (const char* e = std::getenv("B200SIFT_PRIO_A")) {  // development switch: priority of part A relative to the greatest (0, 1, 2 ..: lower)
        prio_a = std::min(prio_least, prio_greatest + std::atoi(e));
    }
    CU(cudaStreamCreateWithFlags(&d->copy_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&d->ev_user, cudaEventDisableTiming));

    for (int si = 0; si < d->nslots; si++) {
        Slot& sl = d->slot[si];
        FrameParams& fp = sl.fp;
        std::memset(&fp, 0, sizeof(fp));
        fp.nOct = d->nOct;
        fp.nScales = nS;
        fp.nLayers = nL;
        fp.o_begin = 0;
        fp.o_end = d->nOct;
        fp.numFeatures = c.numFeatures;
        fp.extrema_thr = (float)std::floor(0.5 * c.contrastThreshould / c.numOctaveLayers * 255 * 1.f);  // Detector.cu:362
        fp.contrast_thr = (float)c.contrastThreshould;
        fp.edge_thr = (float)c.edgeThreshould;
        fp.sigma = (float)c.sigma;
        fp.opt_dedup = d->opt.remove_duplicates ? 1 : 0;
        fp.opt_ori_interp = d->opt.interpolate_orientation ? 1 : 0;
        fp.opt_upscale = d->upscaled ? 1 : 0;

        if (int rc = dev_alloc(d, &sl.d_input, (size_t)d->in_pitch * c.row_width)) return rc;
        if (int rc = dev_alloc(d, &sl.d_input_u8, d->u8_pitch * c.row_width)) return rc;
        if (d->upscaled)
            if (int rc = dev_alloc(d, &sl.d_up, (size_t)d->up_pitch * d->base_h)) return rc;

        int rows_total = 0, scat_ctas = 0;
        for (int o = 0; o < d->nOct; o++) {
            OctParams& oc = fp.oct[o];
            oc.w = (int)roundf(d->base_w / powf(2, o));  // Detector.cu:83 (m_init_col / m_init_row: doubled when upscaled)
            oc.h = (int)roundf(d->base_h / powf(2, o));
            if (oc.w < 1 || oc.h < 1) return fail(B200SIFT_EINVAL, "octave %d is empty", o);
            oc.pitch = pitch_floats(oc.w);
            oc.cap = std::min(oc.pitch * oc.h, c.numFeatures * 10);  // KeyPoint.cu:13, Detector.cu:109
            oc.words = 2 * ((oc.w + 63) / 64);  // two mask words (even / odd pixels) per 64-pixel strip
            oc.row0 = rows_total;
            rows_total += nS * oc.h;
            oc.scat_cta0 = scat_ctas;
            scat_ctas += (nS * oc.h + 7) / 8;
        }
        fp.rows_total = rows_total;
        if (int rc = dev_alloc(d, &sl.d_rowcount_all, (size_t)rows_total)) return rc;
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
            if (int rc = dev_alloc(d, &oc.gauss, ls * nL + oc.pitch)) return rc;
            if (int rc = dev_alloc(d, &oc.mask, (size_t)nS * oc.h * oc.words)) return rc;
            oc.rowcount = sl.d_rowcount_all + oc.row0;
            if (si == 0) {  // candidate lists: ONE set, handed from frame to frame (stale entries and all)
                if (int rc = dev_alloc(d, &oc.cand, (size_t)oc.cap)) return rc;
                if (int rc = dev_alloc(d, &oc.candF, (size_t)oc.cap)) return rc;
                if (int rc = dev_alloc(d, &oc.coll, (size_t)oc.cap)) return rc;
                if (int rc = dev_alloc(d, &oc.collF, (size_t)oc.cap)) return rc;
                if (int rc = dev_alloc(d, &oc.flag, (size_t)oc.cap)) return rc;
                if (int rc = dev_alloc(d, &oc.oricnt, (size_t)oc.cap)) return rc;
            } else {
                const OctParams& o0 = d->slot[0].fp.oct[o];
                oc.cand = o0.cand; oc.candF = o0.candF; oc.coll = o0.coll; oc.collF = o0.collF;
                oc.flag = o0.flag; oc.oricnt = o0.oricnt;
            }
            if (int rc = dev_alloc(d, &oc.fin_kp, (size_t)c.numFeatures)) return rc;
            if (int rc = dev_alloc(d, &oc.fin_ft, (size_t)c.numFeatures)) return rc;
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
        {
            uint8_t* p = nullptr;
            if (int rc = dev_alloc(d, &p, d->match_scratch_bytes_)) return rc;
            sl.match_scratch = p;
            if (int rc = dev_alloc(d, &sl.d_match_idx, (size_t)c.numFeatures)) return rc;
        }
        CU(cudaMallocHost(&sl.h_total, sizeof(int)));
        *sl.h_total = 0;
        CU(cudaStreamCreateWithPriority(&sl.stream, cudaStreamNonBlocking, prio_greatest));
        for (int i = 0; i < kSideStreams; i++) CU(cudaStreamCreateWithPriority(&sl.side[i], cudaStreamNonBlocking, prio_greatest));
        // part A (pyramid + extrema) is captured on streams of their own so that its kernel nodes can carry a lower priority
        // than part B's (the ordered candidate stages, whose latency under load bounds the frame rate)
        CU(cudaStreamCreateWithPriority(&sl.stream_a, cudaStreamNonBlocking, prio_a));
        for (int i = 0; i < kSideStreams; i++) CU(cudaStreamCreateWithPriority(&sl.side_a[i], cudaStreamNonBlocking, prio_a));
        for (int o = 0; o < kMaxOctaves; o++) {
            CU(cudaEventCreateWithFlags(&sl.ev_fork[o], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&sl.ev_join[o], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&sl.ev_counts[o], cudaEventDisableTiming));
        }
        CU(cudaEventCreateWithFlags(&sl.ev_copied, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&sl.ev_consumed, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&sl.ev_done, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&sl.ev_matched, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&sl.ev_chain, cudaEventDisableTiming));
    }

    const FrameParams& fp0 = d->slot[0].fp;
    d->tail_first = d->nOct;
    for (int o = d->nOct - 1; o >= 0 && tail_octave(fp0.oct[o].w, fp0.oct[o].h); o--) d->tail_first = o;
    // graph branches over the multi-CTA octaves [0, tail_first): big octaves alone, the rest together
    d->n_side = 0;
    for (int o = 0; o < d->tail_first;) {
        int e = o + 1;
        const bool big = fp0.oct[o].w * fp0.oct[o].h >= kBranchPixels && d->n_side < kSideStreams - 1;
        if (!big) e = d->tail_first;
        d->br_begin[d->n_side] = o;
        d->br_end[d->n_side] = e;
        d->n_side++;
        o = e;
    }

    // Three graphs per slot (parts A, B, C of the frame).
    d->launches = 0;
    for (int si = 0; si < d->nslots; si++) {
        Slot& sl = d->slot[si];
        const int parts[3] = {kPartA, kPartB, kPartC};
        int total = 0;
        for (int g = 0; g < 3; g++) {
            int n = 0;
            const bool on_a = parts[g] == kPartA;   // part A is captured on its own (possibly lower-priority) streams
            if (on_a) {
                std::swap(sl.stream, sl.stream_a);
                for (int i = 0; i < kSideStreams; i++) std::swap(sl.side[i], sl.side_a[i]);
            }
            cudaError_t e = cudaStreamBeginCapture(sl.stream, cudaStreamCaptureModeThreadLocal);
            int rc = 0;
            if (e == cudaSuccess) {
                rc = enqueue_frame(d, sl, 7, &n, nullptr, parts[g]);
                e = cudaStreamEndCapture(sl.stream, &sl.graph[g]);
            }
            if (on_a) {
                std::swap(sl.stream, sl.stream_a);
                for (int i = 0; i < kSideStreams; i++) std::swap(sl.side[i], sl.side_a[i]);
            }
            if (rc) return rc;
            CU(e);
            CU(cudaGraphInstantiate(&sl.exec[g], sl.graph[g], 0));
            total += n;
        }
        d->launches = total;
        CU(cudaStreamSynchronize(sl.stream));
    }
    return 0;
}

}  // namespace

// The engine keeps up to 32 frames in flight, each on its own stream with up to five graph branches. With the CUDA
// default of 8 hardware work queues per device those streams alias onto the same queues and only ~4 frames really overlap
// (profiles/r2_timeline.md); 32 queues are worth +7.5% end to end at 1920x1200 and +9-11% at 752x480. The variable is read
// when the CUDA context is created, so it is set when the library is loaded (a C++ program that links the library: before
// main), and never over a value the user chose.
__attribute__((constructor)) static void b200sift_on_load() { setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", /*overwrite=*/0); }

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

void b200sift_default_options(b200sift_options* opt) {
    if (!opt) return;
    std::memset(opt, 0, sizeof(*opt));
    opt->frames_in_flight = kDefaultSlots;
}

int b200sift_create(const b200sift_config* cfg, b200sift_detector** out) { return b200sift_create_ex(cfg, nullptr, out); }

int b200sift_create_ex(const b200sift_config* cfg, const b200sift_options* opt, b200sift_detector** out) {
    if (!cfg || !out) return fail(B200SIFT_EINVAL, "null argument");
    *out = nullptr;
    b200sift_options o;
    b200sift_default_options(&o);
    if (opt) o = *opt;
    if (o.frames_in_flight == 0) o.frames_in_flight = kDefaultSlots;
    if (o.frames_in_flight < kMinSlots || o.frames_in_flight > kMaxSlots)
        return fail(B200SIFT_EINVAL, "frames_in_flight must be in %d..%d", kMinSlots, kMaxSlots);
    if (cfg->col_width <= 0 || cfg->row_width <= 0)
        return fail(B200SIFT_EINVAL, "Image width or height not set.");  // Detector.cu:22-25
    if (cfg->upscale && !o.working_upscale)
        return fail(B200SIFT_EINVAL, "upscale=true is broken in the reference (CudaSiftConfig.hh:12): set b200sift_options::working_upscale to opt into the fixed version");
    if (cfg->numOctaveLayers < 1 || cfg->numOctaveLayers > kMaxScales)
        return fail(B200SIFT_EINVAL, "numOctaveLayers must be in 1..%d", kMaxScales);
    if (!(cfg->sigma > 0)) return fail(B200SIFT_EINVAL, "sigma must be positive");
    if (cfg->numFeatures <= 0) return fail(B200SIFT_EINVAL, "numFeatures must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(B200SIFT_ECUDA, "no CUDA device: this engine has no CPU fallback");
    }
    b200sift_detector* d = new b200sift_detector;
    d->cfg = *cfg;
    d->opt = o;
    d->nslots = o.frames_in_flight;
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

static int sync_all(b200sift_detector* d) {
    for (int i = 0; i < d->nslots; i++)
        if (d->slot[i].stream) CU(cudaStreamSynchronize(d->slot[i].stream));
    return 0;
}

int b200sift_destroy(b200sift_detector* d) {
    if (!d) return 0;
    sync_all(d);
    if (d->copy_stream) cudaStreamSynchronize(d->copy_stream);
    for (int si = 0; si < d->nslots; si++) {
        Slot& sl = d->slot[si];
        for (int g = 0; g < 3; g++) {
            if (sl.exec[g]) cudaGraphExecDestroy(sl.exec[g]);
            if (sl.graph[g]) cudaGraphDestroy(sl.graph[g]);
        }
        if (sl.ev_chain) cudaEventDestroy(sl.ev_chain);
        if (sl.ev_copied) cudaEventDestroy(sl.ev_copied);
        if (sl.ev_consumed) cudaEventDestroy(sl.ev_consumed);
        if (sl.ev_done) cudaEventDestroy(sl.ev_done);
        if (sl.ev_matched) cudaEventDestroy(sl.ev_matched);
        if (sl.h_total) cudaFreeHost(sl.h_total);
        if (sl.stream) cudaStreamDestroy(sl.stream);
        if (sl.stream_a) cudaStreamDestroy(sl.stream_a);
        for (int i = 0; i < kSideStreams; i++)
            if (sl.side_a[i]) cudaStreamDestroy(sl.side_a[i]);
        for (int i = 0; i < kSideStreams; i++)
            if (sl.side[i]) cudaStreamDestroy(sl.side[i]);
        for (int o = 0; o < kMaxOctaves; o++) {
            if (sl.ev_fork[o]) cudaEventDestroy(sl.ev_fork[o]);
            if (sl.ev_counts[o]) cudaEventDestroy(sl.ev_counts[o]);
            if (sl.ev_join[o]) cudaEventDestroy(sl.ev_join[o]);
        }
    }
    if (d->copy_stream) cudaStreamDestroy(d->copy_stream);
    if (d->ev_user) cudaEventDestroy(d->ev_user);
    for (void* p : d->allocs) cudaFree(p);
    delete d;
    return 0;
}

int b200sift_num_octaves(const b200sift_detector* d) { return d ? d->nOct : B200SIFT_EINVAL; }

int b200sift_octave_info(const b200sift_detector* d, int o, int* dims) {
    if (!d || !dims || o < 0 || o >= d->nOct) return fail(B200SIFT_EINVAL, "bad octave");
    const OctParams& oc = d->slot[0].fp.oct[o];
    dims[0] = oc.w; dims[1] = oc.h; dims[2] = oc.pitch; dims[3] = oc.cap;
    return 0;
}

// Frame n runs in slot n % nslots. Before the slot is reused, everything that read its previous contents from
// another slot's stream (the matcher / result copies of frame n-2, which used it as "previous frame") must be
// done; host frames are copied on the copy stream, which only waits until the frame that last used the slot
// has consumed its input, so the copy overlaps the frames in flight.
static int begin_frame(b200sift_detector* d, Slot** out, bool on_copy_stream) {
    Slot& sl = d->slot[d->frame_no % d->nslots];
    if (sl.matched_pending) {
        CU(cudaStreamWaitEvent(sl.stream, sl.ev_matched, 0));
        sl.matched_pending = false;
    }
    if (on_copy_stream) CU(cudaStreamWaitEvent(d->copy_stream, sl.ev_consumed, 0));
    *out = &sl;
    return 0;
}
// The three graphs of one frame on the slot's stream. Part B waits for part B of the frame before it
// (`pv`, another slot and stream): the shared candidate lists pass from frame to frame in order.
static int launch_graphs(Slot& sl, Slot* pv) {
    CU(cudaGraphLaunch(sl.exec[0], sl.stream));
    CU(cudaEventRecord(sl.ev_consumed, sl.stream));  // the input image has been read
    if (pv) CU(cudaStreamWaitEvent(sl.stream, pv->ev_chain, 0));
    CU(cudaGraphLaunch(sl.exec[1], sl.stream));
    CU(cudaEventRecord(sl.ev_chain, sl.stream));
    CU(cudaGraphLaunch(sl.exec[2], sl.stream));
    CU(cudaEventRecord(sl.ev_done, sl.stream));
    return 0;
}
static int launch_frame(b200sift_detector* d, Slot& sl, bool from_copy_stream) {
    if (from_copy_stream) {
        CU(cudaEventRecord(sl.ev_copied, d->copy_stream));
        CU(cudaStreamWaitEvent(sl.stream, sl.ev_copied, 0));
    }
    if (int rc = launch_graphs(sl, d->frame_no > 0 ? &d->cur() : nullptr)) return rc;
    d->frame_no++;
    d->have_frame = true;
    return 0;
}

int b200sift_detect_host_async(b200sift_detector* d, const float* img, int rows, int cols) {
    if (!d || !img) return fail(B200SIFT_EINVAL, "null argument");
    if (rows != d->cfg.row_width || cols != d->cfg.col_width)
        return fail(B200SIFT_EINVAL, "image is %dx%d, detector was configured for %dx%d", cols, rows, d->cfg.col_width, d->cfg.row_width);
    Slot* sl = nullptr;
    if (int rc = begin_frame(d, &sl, true)) return rc;
    CU(cudaMemcpy2DAsync(sl->d_input, (size_t)d->in_pitch * 4, img, (size_t)cols * 4, (size_t)cols * 4, rows,
                         cudaMemcpyHostToDevice, d->copy_stream));
    return launch_frame(d, *sl, true);
}

int b200sift_sync(b200sift_detector* d, int* total) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    if (int rc = sync_all(d)) return rc;
    if (total) *total = d->have_frame ? *d->cur().h_total : 0;
    return 0;
}

int b200sift_detect_host(b200sift_detector* d, const float* img, int rows, int cols, int* total) {
    int rc = b200sift_detect_host_async(d, img, rows, cols);
    if (rc) return rc;
    return b200sift_sync(d, total);
}

int b200sift_detect_device_async(b200sift_detector* d, const float* d_img, size_t pitch_bytes, void* producer_stream) {
    if (!d || !d_img) return fail(B200SIFT_EINVAL, "null argument");
    const int cols = d->cfg.col_width, rows = d->cfg.row_width;
    if (pitch_bytes < (size_t)cols * 4) return fail(B200SIFT_EINVAL, "pitch smaller than a row");
    Slot* sl = nullptr;
    if (int rc = begin_frame(d, &sl, false)) return rc;
    // The slot streams are cudaStreamNonBlocking: nothing orders them against the stream that produced the image
    // (not even the legacy default stream), so the copy below waits for an event recorded there.
    CU(cudaEventRecord(d->ev_user, static_cast<cudaStream_t>(producer_stream)));
    CU(cudaStreamWaitEvent(sl->stream, d->ev_user, 0));
    CU(cudaMemcpy2DAsync(sl->d_input, (size_t)d->in_pitch * 4, d_img, pitch_bytes, (size_t)cols * 4, rows,
                         cudaMemcpyDeviceToDevice, sl->stream));
    return launch_frame(d, *sl, false);
}

int b200sift_detect_host_u8_async(b200sift_detector* d, const uint8_t* img, int rows, int cols) {
    if (!d || !img) return fail(B200SIFT_EINVAL, "null argument");
    if (rows != d->cfg.row_width || cols != d->cfg.col_width)
        return fail(B200SIFT_EINVAL, "image is %dx%d, detector was configured for %dx%d", cols, rows, d->cfg.col_width, d->cfg.row_width);
    Slot* sl = nullptr;
    if (int rc = begin_frame(d, &sl, true)) return rc;
    CU(cudaMemcpy2DAsync(sl->d_input_u8, d->u8_pitch, img, (size_t)cols, (size_t)cols, rows, cudaMemcpyHostToDevice, d->copy_stream));
    launch_u8_to_float(sl->d_input_u8, d->u8_pitch, sl->d_input, d->in_pitch, cols, rows, d->copy_stream);
    return launch_frame(d, *sl, true);
}

// Everything enqueued so far on the latest frame's stream has read the previous frame's slot: remember it, so
// that the frame which reuses that slot (two frames from now) waits for these readers.
static int mark_prev_read(b200sift_detector* d) {
    Slot &cur = d->cur(), &prv = d->prev();
    CU(cudaEventRecord(prv.ev_matched, cur.stream));
    prv.matched_pending = true;
    return 0;
}

int b200sift_match_consecutive_async(b200sift_detector* d, int* d_out_idx) {
    if (!d || !d_out_idx) return fail(B200SIFT_EINVAL, "null argument");
    if (d->frame_no == 0) return fail(B200SIFT_ESTATE, "no frame has been processed");
    Slot &cur = d->cur(), &prv = d->prev();
    CU(cudaStreamWaitEvent(cur.stream, prv.ev_done, 0));
    // query = previous frame's descriptors, train = this frame's; both sizes are read on the device
    if (match_launch_counts(prv.fp.out_desc, cur.fp.out_desc, d->cfg.numFeatures, prv.fp.total, cur.fp.total, d_out_idx,
                            cur.match_scratch, d->match_scratch_bytes_, cur.stream))
        return fail(B200SIFT_EINVAL, "matcher scratch too small");
    CU(cudaGetLastError());
    return mark_prev_read(d);
}

int b200sift_copy_results_async(b200sift_detector* d, float* kpts3, float* feats4, uint16_t* desc, int* counts2) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    Slot &cur = d->cur(), &prv = d->prev();
    const size_t n = (size_t)d->cfg.numFeatures;
    if (kpts3) CU(cudaMemcpyAsync(kpts3, cur.fp.out_kpts, n * sizeof(float3), cudaMemcpyDeviceToHost, cur.stream));
    if (feats4) CU(cudaMemcpyAsync(feats4, cur.fp.out_feats, n * sizeof(float4), cudaMemcpyDeviceToHost, cur.stream));
    if (desc) CU(cudaMemcpyAsync(desc, cur.fp.out_desc, n * 128 * sizeof(__half), cudaMemcpyDeviceToHost, cur.stream));
    if (counts2) {
        CU(cudaMemcpyAsync(counts2, cur.fp.total, sizeof(int), cudaMemcpyDeviceToHost, cur.stream));
        CU(cudaStreamWaitEvent(cur.stream, prv.ev_done, 0));
        CU(cudaMemcpyAsync(counts2 + 1, prv.fp.total, sizeof(int), cudaMemcpyDeviceToHost, cur.stream));
        return mark_prev_read(d);
    }
    return 0;
}

int b200sift_detect_batch_host(b200sift_detector* d, const void* const* frames, int n, int rows, int cols, int is_u8,
                               float* const* kpts3, float* const* feats4, uint16_t* const* desc, int* const* counts2,
                               int* const* match_idx) {
    if (!d || !frames || n < 0) return fail(B200SIFT_EINVAL, "bad argument");
    for (int i = 0; i < n; i++) {
        if (!frames[i]) return fail(B200SIFT_EINVAL, "frame %d is null", i);
        int rc = is_u8 ? b200sift_detect_host_u8_async(d, static_cast<const uint8_t*>(frames[i]), rows, cols)
                       : b200sift_detect_host_async(d, static_cast<const float*>(frames[i]), rows, cols);
        if (rc) return rc;
        if (match_idx && match_idx[i]) {
            Slot& cur = d->cur();
            if ((rc = b200sift_match_consecutive_async(d, cur.d_match_idx))) return rc;
            CU(cudaMemcpyAsync(match_idx[i], cur.d_match_idx, (size_t)d->cfg.numFeatures * sizeof(int), cudaMemcpyDeviceToHost, cur.stream));
        }
        if ((rc = b200sift_copy_results_async(d, kpts3 ? kpts3[i] : nullptr, feats4 ? feats4[i] : nullptr, desc ? desc[i] : nullptr,
                                              counts2 ? counts2[i] : nullptr)))
            return rc;
    }
    return sync_all(d);
}

void* b200sift_stream(b200sift_detector* d) { return d ? (void*)d->cur().stream : nullptr; }

int b200sift_results(b200sift_detector* d, const float** kpts, const float** feats, const uint16_t** desc,
                     const uint16_t** prev_desc, int* total) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    const FrameParams& fp = d->cur().fp;
    if (kpts) *kpts = reinterpret_cast<const float*>(fp.out_kpts);
    if (feats) *feats = reinterpret_cast<const float*>(fp.out_feats);
    if (desc) *desc = reinterpret_cast<const uint16_t*>(fp.out_desc);
    if (prev_desc) *prev_desc = reinterpret_cast<const uint16_t*>(d->prev().fp.out_desc);
    if (total) return b200sift_sync(d, total);
    return 0;
}

int b200sift_copy_to_host(b200sift_detector* d, float* kpts3, float* feats4, uint16_t* desc, int with_desc) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    int total = 0;
    int rc = b200sift_sync(d, &total);
    if (rc) return rc;
    Slot& cur = d->cur();
    if (kpts3) CU(cudaMemcpyAsync(kpts3, cur.fp.out_kpts, (size_t)total * sizeof(float3), cudaMemcpyDeviceToHost, cur.stream));
    if (feats4) CU(cudaMemcpyAsync(feats4, cur.fp.out_feats, (size_t)total * sizeof(float4), cudaMemcpyDeviceToHost, cur.stream));
    if (with_desc && desc)
        CU(cudaMemcpyAsync(desc, cur.fp.out_desc, (size_t)total * 128 * sizeof(__half), cudaMemcpyDeviceToHost, cur.stream));
    CU(cudaStreamSynchronize(cur.stream));
    return 0;
}

int b200sift_debug_get_stack(b200sift_detector* d, int o, int which, float* out) {
    if (!d || !out || o < 0 || o >= d->nOct || which != 0) return fail(B200SIFT_EINVAL, "bad argument");
    const OctParams& oc = d->cur().fp.oct[o];
    if (int rc = sync_all(d)) return rc;
    CU(cudaMemcpy2D(out, (size_t)oc.w * 4, oc.gauss, (size_t)oc.pitch * 4, (size_t)oc.w * 4, (size_t)oc.h * d->nLayers,
                    cudaMemcpyDeviceToHost));
    return d->nLayers;
}

int b200sift_debug_get_kpt_buffer(b200sift_detector* d, int o, int which, void* out, int max_elems) {
    if (!d || !out || o < 0 || o >= d->nOct) return fail(B200SIFT_EINVAL, "bad argument");
    const OctParams& oc = d->cur().fp.oct[o];
    const int n = std::min(max_elems, oc.cap);
    if (int rc = sync_all(d)) return rc;
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
    if (int rc = sync_all(d)) return rc;
    int c[4];
    CU(cudaMemcpy(c, d->cur().fp.oct[o].counts, sizeof(c), cudaMemcpyDeviceToHost));
    counts[0] = c[0]; counts[1] = c[1]; counts[2] = c[2];
    return 0;
}

int b200sift_debug_get_taps(b200sift_detector* d, int which, float* out, int max_elems) {
    if (!d || which < -1 || which >= d->nLayers) return fail(B200SIFT_EINVAL, "bad argument");
    const Taps& t = which < 0 ? d->init_taps : d->taps[which];
    if (out && t.n <= max_elems) std::memcpy(out, t.v, t.n * sizeof(float));
    return t.n;
}

// Stage dump of the latest frame (the role of Detector::setDataGen + perf/PerfData.cu:4-85 in the reference,
// without msgpack/zlib): raw little-endian arrays + manifest.json under `dir` (which must exist).
namespace {
bool write_bin(const std::string& path, const void* p, size_t bytes) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = bytes == 0 || std::fwrite(p, 1, bytes, f) == bytes;
    return std::fclose(f) == 0 && ok;
}
}  // namespace

int b200sift_dump_frame(b200sift_detector* d, const char* dir) {
    if (!d || !dir) return fail(B200SIFT_EINVAL, "null argument");
    if (d->frame_no == 0) return fail(B200SIFT_EINVAL, "no frame has been processed");
    if (int rc = sync_all(d)) return rc;
    Slot& sl = d->cur();
    const std::string base = std::string(dir) + "/";
    const int W = d->cfg.col_width, H = d->cfg.row_width;
    std::vector<float> buf((size_t)W * H);
    CU(cudaMemcpy2D(buf.data(), (size_t)W * 4, sl.d_input, (size_t)d->in_pitch * 4, (size_t)W * 4, H, cudaMemcpyDeviceToHost));
    if (!write_bin(base + "input.f32", buf.data(), buf.size() * 4)) return fail(B200SIFT_EINVAL, "cannot write into the dump directory");
    std::string man = "{\n \"format\": \"b200sift-stage-dump-1\",\n";
    man += " \"config\": {\"col_width\": " + std::to_string(W) + ", \"row_width\": " + std::to_string(H) +
           ", \"numFeatures\": " + std::to_string(d->cfg.numFeatures) + ", \"numOctaveLayers\": " + std::to_string(d->cfg.numOctaveLayers) +
           ", \"contrastThreshould\": " + std::to_string(d->cfg.contrastThreshould) + ", \"edgeThreshould\": " + std::to_string(d->cfg.edgeThreshould) +
           ", \"sigma\": " + std::to_string(d->cfg.sigma) + "},\n";
    man += " \"frame_no\": " + std::to_string(d->frame_no) + ",\n \"octaves\": [\n";
    for (int o = 0; o < d->nOct; o++) {
        const OctParams& oc = sl.fp.oct[o];
        const std::string pre = "oct" + std::to_string(o) + "_";
        buf.resize((size_t)oc.w * oc.h * d->nLayers);
        CU(cudaMemcpy2D(buf.data(), (size_t)oc.w * 4, oc.gauss, (size_t)oc.pitch * 4, (size_t)oc.w * 4, (size_t)oc.h * d->nLayers,
                        cudaMemcpyDeviceToHost));
        bool ok = write_bin(base + pre + "gauss.f32", buf.data(), buf.size() * 4);
        int c[4] = {0, 0, 0, 0};
        CU(cudaMemcpy(c, oc.counts, sizeof(c), cudaMemcpyDeviceToHost));
        const int M = std::min(c[1], oc.cap), T = std::min(c[2], oc.cap);
        std::vector<float> k((size_t)oc.cap * 4);
        // after a whole frame: coll[0, M) = refined keypoints in raster order (compact_refined), cand[0, T) = the
        // oriented keypoints in (keypoint, bin) order (compact_oriented); the last entry of each keeps whatever
        // was in its slot before (quirk Q5)
        CU(cudaMemcpy(k.data(), oc.coll, (size_t)M * sizeof(float3), cudaMemcpyDeviceToHost));
        ok = ok && write_bin(base + pre + "refined_kpts.f32x3", k.data(), (size_t)M * sizeof(float3));
        CU(cudaMemcpy(k.data(), oc.collF, (size_t)M * sizeof(float4), cudaMemcpyDeviceToHost));
        ok = ok && write_bin(base + pre + "refined_features.f32x4", k.data(), (size_t)M * sizeof(float4));
        CU(cudaMemcpy(k.data(), oc.cand, (size_t)T * sizeof(float3), cudaMemcpyDeviceToHost));
        ok = ok && write_bin(base + pre + "oriented_kpts.f32x3", k.data(), (size_t)T * sizeof(float3));
        CU(cudaMemcpy(k.data(), oc.candF, (size_t)T * sizeof(float4), cudaMemcpyDeviceToHost));
        ok = ok && write_bin(base + pre + "oriented_features.f32x4", k.data(), (size_t)T * sizeof(float4));
        if (!ok) return fail(B200SIFT_EINVAL, "cannot write into the dump directory");
        man += "  {\"w\": " + std::to_string(oc.w) + ", \"h\": " + std::to_string(oc.h) + ", \"cap\": " + std::to_string(oc.cap) +
               ", \"raw_extrema\": " + std::to_string(c[0]) + ", \"refined\": " + std::to_string(c[1]) + ", \"oriented\": " + std::to_string(c[2]) +
               "}" + (o + 1 < d->nOct ? ",\n" : "\n");
    }
    int total = 0;
    if (d->have_frame) {
        if (int rc = b200sift_sync(d, &total)) return rc;
        std::vector<float> k((size_t)total * 4);
        std::vector<uint16_t> ds((size_t)total * 128);
        CU(cudaMemcpy(k.data(), sl.fp.out_kpts, (size_t)total * sizeof(float3), cudaMemcpyDeviceToHost));
        bool ok = write_bin(base + "final_kpts.f32x3", k.data(), (size_t)total * sizeof(float3));
        CU(cudaMemcpy(k.data(), sl.fp.out_feats, (size_t)total * sizeof(float4), cudaMemcpyDeviceToHost));
        ok = ok && write_bin(base + "final_features.f32x4", k.data(), (size_t)total * sizeof(float4));
        CU(cudaMemcpy(ds.data(), sl.fp.out_desc, ds.size() * 2, cudaMemcpyDeviceToHost));
        ok = ok && write_bin(base + "final_descriptors.f16x128", ds.data(), ds.size() * 2);
        if (!ok) return fail(B200SIFT_EINVAL, "cannot write into the dump directory");
    }
    man += " ],\n \"total_size\": " + std::to_string(total) + "\n}\n";
    if (!write_bin(base + "manifest.json", man.data(), man.size())) return fail(B200SIFT_EINVAL, "cannot write into the dump directory");
    return 0;
}

// Runs the pipeline up to `last_stage` eagerly (no graph) in the next slot, which becomes the latest frame.
int b200sift_debug_run_stages(b200sift_detector* d, const float* img, int rows, int cols, int last_stage) {
    if (!d || !img) return fail(B200SIFT_EINVAL, "null argument");
    if (rows != d->cfg.row_width || cols != d->cfg.col_width) return fail(B200SIFT_EINVAL, "image size mismatch");
    if (int rc = sync_all(d)) return rc;
    Slot& sl = d->slot[d->frame_no % d->nslots];
    sl.matched_pending = false;
    CU(cudaMemcpy2DAsync(sl.d_input, (size_t)d->in_pitch * 4, img, (size_t)cols * 4, (size_t)cols * 4, rows,
                         cudaMemcpyHostToDevice, sl.stream));
    int rc = enqueue_frame(d, sl, last_stage, nullptr, nullptr);
    if (rc) return rc;
    CU(cudaEventRecord(sl.ev_consumed, sl.stream));
    CU(cudaEventRecord(sl.ev_chain, sl.stream));
    CU(cudaEventRecord(sl.ev_done, sl.stream));
    d->frame_no++;
    if (last_stage >= 6) d->have_frame = true;
    return sync_all(d);
}

// Re-runs the pipeline on the latest frame's input up to `last_stage` (the compat shims of the reference's public stage
// methods, Detector.hh:36-43). Eager, blocking; the re-run becomes the latest frame.
int b200sift_rerun_stages(b200sift_detector* d, int last_stage) {
    if (!d) return fail(B200SIFT_EINVAL, "null argument");
    if (d->frame_no == 0) return fail(B200SIFT_ESTATE, "no frame has been processed");
    if (int rc = sync_all(d)) return rc;
    Slot& src = d->cur();
    Slot& sl = d->slot[d->frame_no % d->nslots];
    sl.matched_pending = false;
    CU(cudaMemcpyAsync(sl.d_input, src.d_input, (size_t)d->in_pitch * d->cfg.row_width * sizeof(float), cudaMemcpyDeviceToDevice, sl.stream));
    int rc = enqueue_frame(d, sl, last_stage, nullptr, nullptr);
    if (rc) return rc;
    CU(cudaEventRecord(sl.ev_consumed, sl.stream));
    CU(cudaEventRecord(sl.ev_chain, sl.stream));
    CU(cudaEventRecord(sl.ev_done, sl.stream));
    d->frame_no++;
    if (last_stage >= 6) d->have_frame = true;
    return sync_all(d);
}

int b200sift_launches_per_frame(const b200sift_detector* d) { return d ? d->launches : B200SIFT_EINVAL; }

int b200sift_frames_in_flight(const b200sift_detector* d) { return d ? d->nslots : kDefaultSlots; }

// Latency of one frame graph: the latest frame's graph replayed back to back on its own stream.
int b200sift_time_frames(b200sift_detector* d, int warmup, int iters, float* ms_per_frame) {
    if (!d || !ms_per_frame || iters <= 0) return fail(B200SIFT_EINVAL, "bad argument");
    if (int rc = sync_all(d)) return rc;
    Slot& sl = d->cur();
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    for (int i = 0; i < warmup; i++)
        if (int rc = launch_graphs(sl, nullptr)) return rc;
    CU(cudaStreamSynchronize(sl.stream));
    CU(cudaEventRecord(e0, sl.stream));
    for (int i = 0; i < iters; i++)
        if (int rc = launch_graphs(sl, nullptr)) return rc;
    CU(cudaEventRecord(e1, sl.stream));
    CU(cudaEventSynchronize(e1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    d->have_frame = true;
    *ms_per_frame = ms / (float)iters;
    return 0;
}

// Throughput of the frame graphs with all frame slots in flight: every slot replays its graph on the latest
// frame's input, round robin, each on its own stream.
int b200sift_time_frames_pipelined(b200sift_detector* d, int warmup, int iters, float* ms_per_frame) {
    if (!d || !ms_per_frame || iters <= 0) return fail(B200SIFT_EINVAL, "bad argument");
    if (int rc = sync_all(d)) return rc;
    const int nslots = d->nslots;
    Slot& src = d->cur();
    const size_t in_bytes = (size_t)d->in_pitch * d->cfg.row_width * sizeof(float);
    for (int i = 0; i < nslots; i++)
        if (&d->slot[i] != &src) CU(cudaMemcpy(d->slot[i].d_input, src.d_input, in_bytes, cudaMemcpyDeviceToDevice));
    cudaEvent_t e0, e1[kMaxSlots];
    CU(cudaEventCreate(&e0));
    for (int i = 0; i < nslots; i++) CU(cudaEventCreate(&e1[i]));
    const int first = (int)(d->frame_no % nslots);
    Slot* last = &src;
    for (int i = 0; i < warmup; i++) {
        Slot& sl = d->slot[(first + i) % nslots];
        if (int rc = launch_graphs(sl, last)) return rc;
        last = &sl;
    }
    if (int rc = sync_all(d)) return rc;
    CU(cudaEventRecord(e0, d->slot[0].stream));
    for (int i = 1; i < nslots; i++) CU(cudaStreamWaitEvent(d->slot[i].stream, e0, 0));
    const int total = (iters + nslots - 1) / nslots * nslots;  // a whole number of rotations
    for (int i = 0; i < total; i++) {
        Slot& sl = d->slot[(first + i) % nslots];
        if (int rc = launch_graphs(sl, last)) return rc;
        last = &sl;
    }
    float ms = 0.f;
    for (int i = 0; i < nslots; i++) CU(cudaEventRecord(e1[i], d->slot[i].stream));
    for (int i = 0; i < nslots; i++) {
        float t = 0.f;
        CU(cudaEventSynchronize(e1[i]));
        CU(cudaEventElapsedTime(&t, e0, e1[i]));
        ms = std::max(ms, t);
    }
    cudaEventDestroy(e0);
    for (int i = 0; i < nslots; i++) cudaEventDestroy(e1[i]);
    *ms_per_frame = ms / (float)total;
    return 0;
}

// Development aid: the frame graphs with every slot in flight as b200sift_time_frames_pipelined runs them, with a timing
// event before part A and after parts A, B and C of every frame. ms_out[4 * i + {0, 1, 2, 3}] = the time (ms since the
// common start) at which frame i's stream reached part A / finished A / finished B / finished C.
int b200sift_debug_timeline(b200sift_detector* d, int n_frames, float* ms_out) {
    if (!d || !ms_out || n_frames <= 0 || n_frames > 1024) return fail(B200SIFT_EINVAL, "bad argument");
    if (int rc = sync_all(d)) return rc;
    const int nslots = d->nslots;
    Slot& src = d->cur();
    const size_t in_bytes = (size_t)d->in_pitch * d->cfg.row_width * sizeof(float);
    for (int i = 0; i < nslots; i++)
        if (&d->slot[i] != &src) CU(cudaMemcpy(d->slot[i].d_input, src.d_input, in_bytes, cudaMemcpyDeviceToDevice));
    std::vector<cudaEvent_t> ev((size_t)4 * n_frames);
    for (auto& e : ev) CU(cudaEventCreate(&e));
    cudaEvent_t e0;
    CU(cudaEventCreate(&e0));
    const int first = (int)(d->frame_no % nslots);
    Slot* last = &src;
    for (int i = 0; i < 2 * nslots; i++) {  // warm-up: two rotations
        Slot& sl = d->slot[(first + i) % nslots];
        if (int rc = launch_graphs(sl, last)) return rc;
        last = &sl;
    }
    if (int rc = sync_all(d)) return rc;
    CU(cudaEventRecord(e0, d->slot[0].stream));
    for (int i = 1; i < nslots; i++) CU(cudaStreamWaitEvent(d->slot[i].stream, e0, 0));
    for (int i = 0; i < n_frames; i++) {
        Slot& sl = d->slot[(first + i) % nslots];
        CU(cudaEventRecord(ev[4 * i], sl.stream));
        CU(cudaGraphLaunch(sl.exec[0], sl.stream));
        CU(cudaEventRecord(sl.ev_consumed, sl.stream));
        CU(cudaEventRecord(ev[4 * i + 1], sl.stream));
        CU(cudaStreamWaitEvent(sl.stream, last->ev_chain, 0));
        CU(cudaGraphLaunch(sl.exec[1], sl.stream));
        CU(cudaEventRecord(sl.ev_chain, sl.stream));
        CU(cudaEventRecord(ev[4 * i + 2], sl.stream));
        CU(cudaGraphLaunch(sl.exec[2], sl.stream));
        CU(cudaEventRecord(sl.ev_done, sl.stream));
        CU(cudaEventRecord(ev[4 * i + 3], sl.stream));
        last = &sl;
    }
    if (int rc = sync_all(d)) return rc;
    for (int i = 0; i < 4 * n_frames; i++) CU(cudaEventElapsedTime(&ms_out[i], e0, ev[i]));
    for (auto& e : ev) cudaEventDestroy(e);
    cudaEventDestroy(e0);
    d->have_frame = true;
    return 0;
}

int b200sift_profile_frame(b200sift_detector* d, int warmup, int iters, float* ms_per_class, int* launches_per_class) {
    if (!d || !ms_per_class || iters <= 0) return fail(B200SIFT_EINVAL, "bad argument");
    if (int rc = sync_all(d)) return rc;
    Slot& sl = d->cur();
    Profiler prof;
    prof.s = sl.stream;
    std::vector<double> acc(B200SIFT_K_COUNT, 0.0);
    std::vector<int> cnt(B200SIFT_K_COUNT, 0);
    for (int it = 0; it < warmup + iters; it++) {
        prof.used = 0;
        int rc = enqueue_frame(d, sl, 7, nullptr, &prof);
        if (rc) return rc;
        CU(cudaStreamSynchronize(sl.stream));
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
    if (int rc = sync_all(d)) return rc;
    Slot& sl = d->cur();
    const int stages[5] = {1, 3, 4, 5, 7};
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    for (int i = 0; i < 5; i++) {
        cudaGraph_t g = nullptr;
        cudaGraphExec_t x = nullptr;
        CU(cudaStreamSynchronize(sl.stream));
        CU(cudaStreamBeginCapture(sl.stream, cudaStreamCaptureModeThreadLocal));
        int rc = enqueue_frame(d, sl, stages[i], nullptr, nullptr);
        cudaError_t e = cudaStreamEndCapture(sl.stream, &g);
        if (rc) return rc;
        CU(e);
        CU(cudaGraphInstantiate(&x, g, 0));
        for (int k = 0; k < 5; k++) CU(cudaGraphLaunch(x, sl.stream));
        CU(cudaEventRecord(e0, sl.stream));
        for (int k = 0; k < iters; k++) CU(cudaGraphLaunch(x, sl.stream));
        CU(cudaEventRecord(e1, sl.stream));
        CU(cudaEventSynchronize(e1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        ms5[i] = ms / (float)iters;
        cudaGraphExecDestroy(x);
        cudaGraphDestroy(g);
    }
    // leave complete results behind
    if (int rc = launch_graphs(sl, nullptr)) return rc;
    CU(cudaStreamSynchronize(sl.stream));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return 0;
}

// ---- matcher
namespace {
// One cached scratch buffer per device, shared by every b200sift_match_device call on that device. Calls on
// different streams are serialised through an event recorded after each call's last kernel (the next user of the
// scratch, on whatever stream, waits for it), so concurrent matchers cannot overwrite each other's survivor records.
std::mutex g_scratch_mu;
void* g_scratch[64] = {nullptr};
size_t g_scratch_bytes[64] = {0};
cudaEvent_t g_scratch_ev[64] = {nullptr};
cudaStream_t g_scratch_last[64] = {nullptr};
bool g_scratch_used[64] = {false};
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
        if (!g_scratch_ev[dev]) CU(cudaEventCreateWithFlags(&g_scratch_ev[dev], cudaEventDisableTiming));
        if (g_scratch_used[dev] && g_scratch_last[dev] != s) CU(cudaStreamWaitEvent(s, g_scratch_ev[dev], 0));
        if (int rc = match_launch(reinterpret_cast<const __half*>(d_query), nq, reinterpret_cast<const __half*>(d_train), nt, d_out_idx,
                                  scratch, need, s))
            return fail(B200SIFT_EINVAL, rc == -2 ? "train descriptors: tensor map rejected (the pointer must be 16-byte aligned)" : "matcher scratch too small");
        CU(cudaGetLastError());
        CU(cudaEventRecord(g_scratch_ev[dev], s));
        g_scratch_last[dev] = s;
        g_scratch_used[dev] = true;
    }
    return 0;
}

// Host-buffer variant (the blocking call sift_cuda::matchBruteForce makes for host data): per-device staging
// buffers (device + pinned host) grow on demand and are kept, so a call costs two H2D copies, the matcher and one
// D2H copy, no allocation. Calls are serialised per process (the reference's matcher is not thread-safe either).
namespace {
struct MatchStage {
    uint16_t *dq = nullptr, *dt = nullptr, *hq = nullptr, *ht = nullptr;
    int *dout = nullptr, *hout = nullptr;
    size_t cq = 0, ct = 0;
    cudaStream_t s = nullptr;
};
std::mutex g_stage_mu;
MatchStage g_stage[64];

int grow_stage(uint16_t** d, uint16_t** h, size_t* cap, size_t rows, int** dout, int** hout) {
    if (*cap >= rows) return 0;
    const size_t n = std::max<size_t>(rows + rows / 2, 256);
    if (*d) { CU(cudaFree(*d)); *d = nullptr; }
    if (*h) { CU(cudaFreeHost(*h)); *h = nullptr; }
    *cap = 0;
    CU(cudaMalloc(d, n * 256));
    CU(cudaMallocHost(h, n * 256));
    if (dout) {
        if (*dout) { CU(cudaFree(*dout)); *dout = nullptr; }
        if (*hout) { CU(cudaFreeHost(*hout)); *hout = nullptr; }
        CU(cudaMalloc(dout, n * sizeof(int)));
        CU(cudaMallocHost(hout, n * sizeof(int)));
    }
    *cap = n;
    return 0;
}
}  // namespace

int b200sift_match_host(const uint16_t* query, int nq, const uint16_t* train, int nt, int* out_idx) {
    if (nq < 0 || nt < 0 || (nq > 0 && (!query || !out_idx)) || (nt > 0 && !train)) return fail(B200SIFT_EINVAL, "bad argument");
    if (nq == 0) return 0;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(B200SIFT_EINVAL, "device index out of range");
    std::lock_guard<std::mutex> lk(g_stage_mu);
    MatchStage& st = g_stage[dev];
    if (!st.s) CU(cudaStreamCreateWithFlags(&st.s, cudaStreamNonBlocking));
    if (int rc = grow_stage(&st.dq, &st.hq, &st.cq, (size_t)nq, &st.dout, &st.hout)) return rc;
    if (int rc = grow_stage(&st.dt, &st.ht, &st.ct, std::max<size_t>((size_t)nt, 1), nullptr, nullptr)) return rc;
    // pageable -> pinned on the host, then asynchronous copies: the query copy overlaps the staging of the train set
    std::memcpy(st.hq, query, (size_t)nq * 256);
    CU(cudaMemcpyAsync(st.dq, st.hq, (size_t)nq * 256, cudaMemcpyHostToDevice, st.s));
    if (nt) {
        std::memcpy(st.ht, train, (size_t)nt * 256);
        CU(cudaMemcpyAsync(st.dt, st.ht, (size_t)nt * 256, cudaMemcpyHostToDevice, st.s));
    }
    if (int rc = b200sift_match_device(st.dq, nq, st.dt, nt, st.dout, st.s)) return rc;
    CU(cudaMemcpyAsync(st.hout, st.dout, (size_t)nq * sizeof(int), cudaMemcpyDeviceToHost, st.s));
    CU(cudaStreamSynchronize(st.s));
    std::memcpy(out_idx, st.hout, (size_t)nq * sizeof(int));
    return 0;
}

int b200sift_match_launches(int nq, int nt) { return match_num_launches(nq, nt); }

int b200sift_match_set_path(int path) {
    if (path < 0 || path > 3)
        return fail(B200SIFT_EINVAL, "path must be 0 (tensor core), 1 (CUDA core), 2 (tensor core, unpipelined large-set filter) or 3 (tensor core, single-pass streaming filter)");
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
