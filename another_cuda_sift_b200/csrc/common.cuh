// Shared device-side structures and stage launchers of the B200 SIFT engine.
//
// Data layout in HBM (one slab per detector, allocated once):
//   * input image            : pitched float, pitch = round_up(w*4, 512) bytes
//   * Gaussian stack/octave  : numOctaveLayers + 3 layers, layer-major then row-major, row pitch rounded up to
//                              512 B, layer stride pitch*h floats -- the layout the reference gets
//                              from cudaMallocPitch (types/CudaImage.cu:192-207), kept because its
//                              descriptor kernel reads one element past a row / layer (quirk Q12)
//   * no DoG stack           : DoG[l] = G[l+1]-G[l] is one exact FADD, recomputed where needed
//   * extrema bitmask        : 1 bit per pixel per scale (numOctaveLayers scales), + per-row popcounts
//   * keypoint slots/octave  : cand/candF/coll/collF with the reference's slot semantics
//                              (types/KeyPoint.cuh:9-29)
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200sift {

constexpr int kMaxOctaves = 16;
// numOctaveLayers (n) and sigma are runtime parameters, as in the reference (interface/Detector.cu:63-94: n + 3 Gaussian
// layers, n + 2 DoG layers, n scales per octave; utils/GaussianUtils.cc:39-67: any tap count). Bounds of this engine:
constexpr int kMaxScales = 6;                  // numOctaveLayers <= 6
constexpr int kMaxLayers = kMaxScales + 3;
constexpr int kBorder = 5;        // SIFT_IMG_BORDER (sift_func/SiftOps.cuh:8)
constexpr int kOriBins = 36;      // sift_func/SiftOps.cu:261
constexpr int kMaxTaps = 85;      // blur radius <= 42, i.e. layer sigma <= ~13.9 (numOctaveLayers = 1 at sigma 1.6 needs 67 taps)
constexpr int kTailPixels = 4096;  // octaves up to this many pixels (and kTailMaxDim per side) go to the single-CTA tail kernel
constexpr int kTailMaxDim = 1024;
inline bool tail_octave(int w, int h) { return w * h <= kTailPixels && w <= kTailMaxDim && h <= kTailMaxDim; }

struct Taps {
    float v[kMaxTaps];
    int n;
};

struct OctParams {
    float* gauss;        // [nLayers][h][pitch]
    uint32_t* mask;      // [nScales][h][words] extrema bits
    int* rowcount;       // [nScales*h] extrema per (scale,row); zeroed every frame
    float3* cand;        // [cap]
    float4* candF;       // [cap]
    float3* coll;        // [cap]
    float4* collF;       // [cap]
    float3* fin_kp;      // [numFeatures] final keypoints of the octave (per frame slot; cand.. are shared by the slots)
    float4* fin_ft;      // [numFeatures]
    int* flag;           // [cap] refine accept flags
    int* oricnt;         // [cap] peaks emitted per refined keypoint
    int* counts;         // [4]: P (raw extrema), M (refined), T (oriented), n = min(P, cap)
    int w, h, pitch, cap;
    int words;           // mask words per row = 2*ceil(w/64)
    int det_task0;       // first warp task of this octave in the detect kernel
    int det_cols, det_rows;  // column strips (64 px) x row strips (kDetRows rows)
    int row0;            // first (scale,row) index of this octave in the scatter kernel
    int scat_cta0;       // first CTA of this octave in the scatter kernel (8 rows per CTA, CTAs never straddle octaves)
};

struct FrameParams {
    OctParams oct[kMaxOctaves];
    int nOct;
    int nScales, nLayers;  // numOctaveLayers, numOctaveLayers + 3
    int o_begin, o_end;  // octave range this launch works on (the frame graph runs octave 0 and the rest as parallel branches)
    int numFeatures;
    int det_tasks;       // total warp tasks of the detect kernel
    int rows_total;      // total (scale,row) entries over all octaves
    float3* out_kpts;    // [numFeatures]
    float4* out_feats;   // [numFeatures]
    __half* out_desc;    // [numFeatures*128]
    int* total;          // device: keypoints returned by this slot's frame
    float extrema_thr;   // floor(0.5*contrast/nLayers*255) (interface/Detector.cu:362)
    float contrast_thr, edge_thr, sigma;
    // opt-in fixes (b200sift_options; all 0 = the reference's behaviour)
    int opt_dedup, opt_ori_interp, opt_upscale;
};

constexpr int kDetRows = 16;  // rows per warp task in the detect kernel

// ---- stage launchers (each enqueues on `s`; no host synchronisation inside) ----
void launch_u8_to_float(const uint8_t* in, size_t in_pitch, float* out, int out_pitch, int w, int h, cudaStream_t s);
// Pyramid launchers (pyramid.cu). `chain`: up to three successive blurs in ONE launch (2-D tiles with recomputed
// halos), optionally fed by the bilinear 2x resample of the previous octave; every stage's layer is stored.
struct BlurStage {
    float* out;     // the layer this stage writes (same pitch / dims as the octave)
    const Taps* taps;
};
bool blur_chain_supported(bool from_prev_octave, const BlurStage* st, int n_stages);
cudaError_t pyramid_setup();  // once per device, outside stream capture (opt-in shared-memory sizes)
bool pyramid_failed();        // true (once) if a launcher could not encode its TMA tensor map since the last call
// src: the layer to blur (pitch src_pitch); or, with from_prev_octave, layer n of the previous octave (src_w x src_h),
// resampled on the fly, the resampled base layer also being written to `base_out`.
void launch_blur_chain(const float* src, int src_pitch, int src_w, int src_h, bool from_prev_octave, float* base_out,
                       const BlurStage* st, int n_stages, int pitch, int w, int h, cudaStream_t s);
void launch_blur(const float* in, int in_pitch, float* out, int out_pitch, int w, int h, const Taps& taps, cudaStream_t s);
void launch_downsample(const float* in, int in_pitch, int iw, int ih, float* out, int out_pitch, int ow, int oh, cudaStream_t s);
// Single-CTA walk over the octaves [first_octave, nOct) (all of at most kTailPixels pixels): layers [l_begin, l_end)
// of each; with l_begin == 0 the base layer comes from the previous octave (or, for octave 0, from `input` blurred
// with `init_taps`), otherwise layer l_begin - 1 is read back from the stack.
void launch_pyramid_tail(const FrameParams& fp, const Taps* taps, const Taps& init_taps, const float* input, int in_pitch,
                         int first_octave, int l_begin, int l_end, cudaStream_t s);
void launch_detect(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s);       // bitmask + row counts
void launch_scatter(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s);      // row offsets, P, ordered candidate list
void launch_refine(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s);       // Taylor refinement + tests
void launch_compact_refined(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s);
void launch_orientation(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s);
void launch_compact_oriented(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s);
void launch_dedup(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s);            // opt-in: two launches
cudaError_t descriptor_setup();
cudaError_t launch_descriptor(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s);

int match_launch(const __half* q, int nq, const __half* t, int nt, int* out, void* scratch, size_t scratch_bytes, cudaStream_t s,
                 float* dbg = nullptr, cudaEvent_t* marks = nullptr);
int match_launch_counts(const __half* q, const __half* t, int max_n, const int* cnt_q, const int* cnt_t, int* out, void* scratch,
                        size_t scratch_bytes, cudaStream_t s);
int match_launch_simt(const __half* q, int nq, const __half* t, int nt, int* out, void* scratch, size_t scratch_bytes, cudaStream_t s);
void match_set_path(int path);
int match_get_path();
size_t match_scratch_bytes(int nq, int nt);
int match_num_launches(int nq, int nt);

inline FrameParams octave_range(const FrameParams& fp, int o_begin, int o_end) {
    FrameParams r = fp;
    r.o_begin = o_begin;
    r.o_end = o_end;
    return r;
}

// reflect-101 border index (same mapping as image_func/Filter.cuh:52-66).
__host__ __device__ inline int reflect101(int idx, int length) {
    if (length <= 1) return 0;
    idx = idx < 0 ? -idx : idx;
    if (idx >= length) {
        int period = 2 * (length - 1);
        idx = idx % period;
        if (idx >= length) idx = period - idx;
    }
    return idx;
}

#ifdef __CUDACC__
// Exclusive scan over n small integers by MANY independent CTAs of kScanBlock threads with one barrier and no
// inter-CTA communication: CTA `seg` owns elements [seg*kScanBlock, (seg+1)*kScanBlock) and recomputes, from
// the whole input (n is a few thousand, L2-resident), the sum of everything before its segment and the
// grand total. emit(i, v, exclusive_prefix, total) is called for the CTA's own elements. Returns the total.
// Small CTAs on purpose: they have to find room next to the descriptor CTAs of the other graph branches.
constexpr int kScanBlock = 256;
template <class Load, class Emit>
__device__ __forceinline__ int segment_scan(int n, int seg, Load load, Emit emit) {
    constexpr int NW = kScanBlock / 32;
    __shared__ int s_before[NW], s_all[NW], s_mine[NW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seg0 = seg * kScanBlock;
    int before = 0, all = 0, mine = 0;
    for (int i = tid; i < n; i += kScanBlock) {
        const int v = load(i);
        all += v;
        if (i < seg0) before += v;
        if (i == seg0 + tid) mine = v;
    }
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += u;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        before += __shfl_xor_sync(0xffffffffu, before, d);
        all += __shfl_xor_sync(0xffffffffu, all, d);
    }
    if (lane == 31) s_mine[warp] = incl;
    if (lane == 0) { s_before[warp] = before; s_all[warp] = all; }
    __syncthreads();
    int base = lane < NW ? s_before[lane] : 0, total = lane < NW ? s_all[lane] : 0;
    int woff = lane < warp ? s_mine[lane] : 0;  // warp < NW
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        base += __shfl_xor_sync(0xffffffffu, base, d);
        total += __shfl_xor_sync(0xffffffffu, total, d);
        woff += __shfl_xor_sync(0xffffffffu, woff, d);
    }
    const int i = seg0 + tid;
    if (i < n) emit(i, mine, base + woff + incl - mine, total);
    return total;
}
#endif

}  // namespace b200sift
