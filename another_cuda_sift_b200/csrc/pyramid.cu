// Gaussian scale-space kernels: fused separable blur (vertical then horizontal in ONE launch,
// the intermediate never leaves shared memory) and the bilinear 2x downsample.
//
// Replaces image_func/Filter.cu:8-51 + FilterImpl.cuh:6-45 (two global-memory passes per layer)
// and image_func/Resize.cu:6-64. Arithmetic is bit-identical to the reference: every output is
// the same in-order chain acc = fmaf(pixel, tap, acc) over offsets -r..+r starting from 0, first
// over rows, then over columns of the row-filtered image, with reflect-101 indices at the image
// border in both passes.
//
// Roofline: HBM-bound. Algorithmic bytes per layer = 8*w*h (read once, write once).
#include "common.cuh"

namespace b200sift {

namespace {

constexpr int kTW = 128;      // output columns per CTA (one warp row = 32 lanes x 4 outputs)
constexpr int kRB = 16;       // output rows per vertical register block
constexpr int kBlurThreads = 320;  // 10 warps: one vertical-pass sweep covers the 128 + 2R columns of two row blocks

// One CTA: TH = nb*kRB output rows x kTW output columns.
//  pass 1 (vertical):   each thread owns one column of one row block, keeps kRB+2R input pixels
//                       in registers (all loads issued up front) and emits kRB row-filtered
//                       values into shared memory. Columns cover the tile plus an R-wide halo,
//                       taken at reflect-101 positions, because the horizontal pass of the
//                       reference reads the row-filtered image at reflected columns.
//  pass 2 (horizontal): each lane produces 4 consecutive outputs from 4+2R shared-memory values
//                       fetched as conflict-free LDS.128, stores one STG.128.
template <int R>
__global__ void __launch_bounds__(kBlurThreads)
blur_vh_kernel(const float* __restrict__ in, int in_pitch, float* __restrict__ out, int out_pitch,
               int w, int h, const __grid_constant__ Taps taps, int nb) {
    constexpr int RP = (R + 3) & ~3;       // halo rounded up so that rows stay 16-byte aligned
    constexpr int TWP = kTW + 2 * RP;      // shared row length
    constexpr int NC = kTW + 2 * R;        // columns that are really needed
    constexpr int NT = 2 * R + 1;
    extern __shared__ __align__(16) float tmp[];  // [nb*kRB][TWP]

    const int x0 = blockIdx.x * kTW;
    const int y0 = blockIdx.y * nb * kRB;
    const int tid = threadIdx.x;

    for (int t = tid; t < NC * nb; t += kBlurThreads) {
        const int b = t / NC;
        const int c = (RP - R) + (t - b * NC);
        const int x = reflect101(x0 - RP + c, w);
        const int ybase = y0 + b * kRB - R;
        float v[kRB + 2 * R];
        if (ybase >= 0 && ybase + kRB + 2 * R <= h) {
            // one 64-bit pointer stepped by the row pitch: a single IMAD.WIDE per load instead of the four address
            // instructions (IADD3 / IMAD.X / LEA / LEA.HI.X) an element-indexed p[i * pitch] compiles to
            const float* p = in + (size_t)ybase * in_pitch + x;
#pragma unroll
            for (int i = 0; i < kRB + 2 * R; i++) {
                v[i] = __ldg(p);
                asm volatile("mad.wide.s32 %0, %1, 4, %0;" : "+l"(p) : "r"(in_pitch));
            }
        } else {
#pragma unroll
            for (int i = 0; i < kRB + 2 * R; i++) v[i] = __ldg(in + (size_t)reflect101(ybase + i, h) * in_pitch + x);
        }
        float* dst = tmp + (size_t)(b * kRB) * TWP + c;
#pragma unroll
        for (int r = 0; r < kRB; r++) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < NT; k++) acc = __fmaf_rn(v[r + k], taps.v[k], acc);
            dst[r * TWP] = acc;
        }
    }
    __syncthreads();

    const int warp = tid >> 5, lane = tid & 31;
    const int TH = nb * kRB;
    const int xo = 4 * lane;
    const int x = x0 + xo;
    if (x >= w) return;
    for (int row = warp; row < TH; row += kBlurThreads / 32) {
        const int y = y0 + row;
        if (y >= h) break;
        constexpr int NV = (2 * RP + 4) / 4;
        float q[NV * 4];
        const float4* src = reinterpret_cast<const float4*>(tmp + (size_t)row * TWP + xo);
#pragma unroll
        for (int m = 0; m < NV; m++) {
            float4 f = src[m];
            q[4 * m] = f.x; q[4 * m + 1] = f.y; q[4 * m + 2] = f.z; q[4 * m + 3] = f.w;
        }
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < NT; k++) acc = __fmaf_rn(q[RP - R + j + k], taps.v[k], acc);
            o[j] = acc;
        }
        float* dst = out + (size_t)y * out_pitch + x;
        if (x + 3 < w) {
            *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
        } else {
            for (int j = 0; j < 4 && x + j < w; j++) dst[j] = o[j];
        }
    }
}

template <int R>
void launch_blur_r(const float* in, int in_pitch, float* out, int out_pitch, int w, int h, const Taps& taps, cudaStream_t s) {
    constexpr int RP = (R + 3) & ~3;
    constexpr int TWP = kTW + 2 * RP;
    // Taller tiles amortise the vertical halo on big images; short tiles spread small images
    // over more SMs.
    const int gx = (w + kTW - 1) / kTW;
    int nb = 4;
    while (nb > 1 && gx * ((h + nb * kRB - 1) / (nb * kRB)) < 2 * 148) nb >>= 1;
    dim3 grid(gx, (h + nb * kRB - 1) / (nb * kRB));
    size_t smem = (size_t)nb * kRB * TWP * sizeof(float);
    blur_vh_kernel<R><<<grid, kBlurThreads, smem, s>>>(in, in_pitch, out, out_pitch, w, h, taps, nb);
}

// Bilinear resample with the reference's exact expression tree (image_func/Resize.cu:26-63);
// the tail is the contraction ptxas chose for the reference on sm_100a:
//   out = fma(fma(v11, 1-xf, v21*xf), 1-yf, fma(v12, 1-xf, v22*xf) * yf)
__device__ __forceinline__ float bilinear_sample(const float* __restrict__ in, int in_pitch, int iw, int ih, int ow, int oh, int x, int y) {
    const float fiw = (float)iw, fih = (float)ih, fow = (float)ow, foh = (float)oh;
    const float coord_x = __fsub_rn(__fdiv_rn(__fmul_rn(__fadd_rn((float)x, 0.5f), fiw), fow), 0.5f);
    const float coord_y = __fsub_rn(__fdiv_rn(__fmul_rn(__fadd_rn((float)y, 0.5f), fih), foh), 0.5f);
    const int max_x = iw - 1, max_y = ih - 1;
    int x1 = min(max_x, max(0, __float2int_rd(coord_x)));
    int y1 = min(max_y, max(0, __float2int_rd(coord_y)));
    const float xf = __fsub_rn(coord_x, truncf(coord_x));
    const float yf = __fsub_rn(coord_y, truncf(coord_y));
    int x2 = xf == 0.f ? x1 + 1 : __float2int_ru(coord_x);
    x2 = min(max_x, max(0, x2));
    int y2 = yf == 0.f ? y1 + 1 : __float2int_ru(coord_y);
    y2 = min(max_y, max(0, y2));
    const float x2e = __fsub_rn(1.f, xf), y2e = __fsub_rn(1.f, yf);
    const float v11 = in[(size_t)y1 * in_pitch + x1], v21 = in[(size_t)y1 * in_pitch + x2];
    const float v12 = in[(size_t)y2 * in_pitch + x1], v22 = in[(size_t)y2 * in_pitch + x2];
    const float a = __fmaf_rn(v11, x2e, __fmul_rn(v21, xf));
    const float r12 = __fmul_rn(__fmaf_rn(v12, x2e, __fmul_rn(v22, xf)), yf);
    return __fmaf_rn(a, y2e, r12);
}

__global__ void __launch_bounds__(256)
downsample_kernel(const float* __restrict__ in, int in_pitch, int iw, int ih,
                  float* __restrict__ out, int out_pitch, int ow, int oh) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31);
    const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= ow || y >= oh) return;
    out[(size_t)y * out_pitch + x] = bilinear_sample(in, in_pitch, iw, ih, ow, oh, x, y);
}

// ---------------------------------------------------------------- pyramid tail
// Octaves of at most kTailPixels pixels are pure launch latency as separate kernels (5 blurs + 1 resample
// each, a few microseconds of dependent launches per layer). ONE CTA walks all of them instead: the
// current layer and the row-filtered intermediate live in shared memory, every layer is also written to
// its place in the Gaussian stack. Arithmetic is the same in-order fmaf chain as blur_vh_kernel.
struct TailParams {
    Taps taps[kGaussLayers];
    float* gauss[kMaxOctaves];
    int w[kMaxOctaves], h[kMaxOctaves], pitch[kMaxOctaves];
    int first, nOct;
};

constexpr int kTailThreads = 1024;
constexpr int kTailR = 10;  // largest blur radius (21 taps)

// One blur layer of the tail: vertical pass s_cur -> s_tmp, horizontal pass s_tmp -> s_cur (+ global).
template <int R>
__device__ __forceinline__ void tail_layer(const Taps& t, float* s_cur, float* s_tmp, const int* s_rtab, const int* s_ctab,
                                           float* __restrict__ dst, int w, int pitch, int n, int tid) {
    constexpr int NT = 2 * R + 1;
    float tap[NT];
#pragma unroll
    for (int k = 0; k < NT; k++) tap[k] = t.v[k];
    // i / w without an integer division: n <= 4096, so the float floor of (i + 0.5) / w is exact
    const float inv_w = __frcp_rn((float)w);
    for (int i = tid; i < n; i += kTailThreads) {
        const int y = __float2int_rd(((float)i + 0.5f) * inv_w), x = i - y * w;
        const int* rt = s_rtab + (y - R + kTailR);
        float v[NT];
#pragma unroll
        for (int k = 0; k < NT; k++) v[k] = s_cur[rt[k] + x];
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < NT; k++) acc = __fmaf_rn(v[k], tap[k], acc);
        s_tmp[i] = acc;
    }
    __syncthreads();
    for (int i = tid; i < n; i += kTailThreads) {
        const int y = __float2int_rd(((float)i + 0.5f) * inv_w), x = i - y * w;
        const int* ct = s_ctab + (x - R + kTailR);
        const float* row = s_tmp + y * w;
        float v[NT];
#pragma unroll
        for (int k = 0; k < NT; k++) v[k] = row[ct[k]];
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < NT; k++) acc = __fmaf_rn(v[k], tap[k], acc);
        s_cur[i] = acc;
        dst[(size_t)y * pitch + x] = acc;
    }
    __syncthreads();  // also orders this CTA's global writes of layer 3 before the next octave reads them
}

__global__ void __launch_bounds__(kTailThreads)
pyramid_tail_kernel(const __grid_constant__ TailParams tp) {
    __shared__ float s_cur[kTailPixels];
    __shared__ float s_tmp[kTailPixels];
    // reflect-101 lookup tables for offsets -kTailR .. dim+kTailR (row table pre-multiplied by the width):
    // the per-tap index is one shared-memory read instead of a modulo
    __shared__ int s_rtab[kTailMaxDim + 2 * kTailR];
    __shared__ int s_ctab[kTailMaxDim + 2 * kTailR];
    const int tid = threadIdx.x;
    for (int o = tp.first; o < tp.nOct; o++) {
        const int w = tp.w[o], h = tp.h[o], pitch = tp.pitch[o], n = w * h;
        float* g = tp.gauss[o];
        const size_t ls = (size_t)pitch * h;
        for (int j = tid; j < h + 2 * kTailR; j += kTailThreads) s_rtab[j] = reflect101(j - kTailR, h) * w;
        for (int j = tid; j < w + 2 * kTailR; j += kTailThreads) s_ctab[j] = reflect101(j - kTailR, w);
        if (o == tp.first) {  // base layer was produced by the launch before this kernel
            for (int i = tid; i < n; i += kTailThreads) s_cur[i] = g[(size_t)(i / w) * pitch + i % w];
        } else {              // bilinear resample of the previous octave's layer 3 (written by this CTA)
            const float* src = tp.gauss[o - 1] + (size_t)tp.pitch[o - 1] * tp.h[o - 1] * kScales;
            for (int i = tid; i < n; i += kTailThreads) {
                const int y = i / w, x = i - y * w;
                const float v = bilinear_sample(src, tp.pitch[o - 1], tp.w[o - 1], tp.h[o - 1], w, h, x, y);
                s_cur[i] = v;
                g[(size_t)y * pitch + x] = v;
            }
        }
        __syncthreads();
        for (int l = 1; l < kGaussLayers; l++) {
            float* dst = g + ls * l;
            switch (tp.taps[l].n / 2) {  // compile-time radius: both passes fully unrolled
                case 4: tail_layer<4>(tp.taps[l], s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
                case 5: tail_layer<5>(tp.taps[l], s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
                case 6: tail_layer<6>(tp.taps[l], s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
                case 8: tail_layer<8>(tp.taps[l], s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
                default: tail_layer<10>(tp.taps[l], s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
            }
        }
    }
}

__global__ void __launch_bounds__(256)
u8_to_float_kernel(const uint8_t* __restrict__ in, size_t in_pitch, float* __restrict__ out, int out_pitch, int w, int h) {
    const int x = (blockIdx.x * 256 + threadIdx.x) * 4;
    const int y = blockIdx.y;
    if (x >= w) return;
    const uint8_t* p = in + (size_t)y * in_pitch + x;
    float* q = out + (size_t)y * out_pitch + x;
    if (x + 3 < w && ((in_pitch & 3) == 0)) {
        uchar4 u = *reinterpret_cast<const uchar4*>(p);
        *reinterpret_cast<float4*>(q) = make_float4((float)u.x, (float)u.y, (float)u.z, (float)u.w);
    } else {
        for (int j = 0; j < 4 && x + j < w; j++) q[j] = (float)p[j];
    }
}

}  // namespace

void launch_blur(const float* in, int in_pitch, float* out, int out_pitch, int w, int h, const Taps& taps, cudaStream_t s) {
    switch (taps.n / 2) {
        case 4: launch_blur_r<4>(in, in_pitch, out, out_pitch, w, h, taps, s); break;
        case 5: launch_blur_r<5>(in, in_pitch, out, out_pitch, w, h, taps, s); break;
        case 6: launch_blur_r<6>(in, in_pitch, out, out_pitch, w, h, taps, s); break;
        case 8: launch_blur_r<8>(in, in_pitch, out, out_pitch, w, h, taps, s); break;
        case 10: launch_blur_r<10>(in, in_pitch, out, out_pitch, w, h, taps, s); break;
        default: break;  // engine.cu rejects other sigmas at create time
    }
}

void launch_downsample(const float* in, int in_pitch, int iw, int ih, float* out, int out_pitch, int ow, int oh, cudaStream_t s) {
    dim3 grid((ow + 31) / 32, (oh + 7) / 8);
    downsample_kernel<<<grid, 256, 0, s>>>(in, in_pitch, iw, ih, out, out_pitch, ow, oh);
}

void launch_pyramid_tail(const FrameParams& fp, const Taps* taps, int first, cudaStream_t s) {
    TailParams tp;
    for (int l = 0; l < kGaussLayers; l++) tp.taps[l] = taps[l];
    for (int o = 0; o < fp.nOct; o++) {
        tp.gauss[o] = fp.oct[o].gauss;
        tp.w[o] = fp.oct[o].w; tp.h[o] = fp.oct[o].h; tp.pitch[o] = fp.oct[o].pitch;
    }
    tp.first = first;
    tp.nOct = fp.nOct;
    pyramid_tail_kernel<<<1, kTailThreads, 0, s>>>(tp);
}

void launch_u8_to_float(const uint8_t* in, size_t in_pitch, float* out, int out_pitch, int w, int h, cudaStream_t s) {
    dim3 grid((w + 1023) / 1024, h);
    u8_to_float_kernel<<<grid, 256, 0, s>>>(in, in_pitch, out, out_pitch, w, h);
}

}  // namespace b200sift
