// Gaussian scale-space kernels.
//
// Replaces image_func/Filter.cu:8-51 + FilterImpl.cuh:6-45 (two global-memory passes per layer),
// image_func/Resize.cu:6-64 and the per-layer launch loop of interface/Detector.cu:274-305.
//
// Arithmetic is bit-identical to the reference: every blurred value is the same in-order chain
// acc = fmaf(pixel, tap, acc) over offsets -r..+r starting from 0, first over rows, then over columns of
// the row-filtered image, with reflect-101 indices at the image border in both passes; the 2x resample is
// the reference's expression tree with the FMA contraction of its sm_100a SASS.
//
//  blur_chain_kernel   K <= 3 successive blurs of one octave in ONE launch, 2-D tiles: the CTA computes stage k
//                      on its output tile grown by the radii of the later stages (the halo is recomputed, never
//                      exchanged), every intermediate lives in shared memory, every stage's tile goes to its
//                      layer of the Gaussian stack. Stage 0 is either a plain layer or the bilinear 2x resample
//                      of the previous octave's layer n, computed on the fly (and stored as the base layer).
//                      The frame graph runs one such launch per octave on the critical path
//                      (resample + L1..Ln) and one on a side branch (L(n+1), L(n+2)), 64 x 32 tiles.
//  blur_vh_kernel      one layer per launch (octave 0: its 2.3 Mpx are throughput-, not latency-bound; any unfused
//                      radius <= 16): vertical pass from global memory in registers, horizontal pass from shared memory.
//  blur_generic_kernel any radius <= 42 at run time (uncommon sigma / numOctaveLayers), one stage.
//  pyramid_tail_kernel ONE CTA walks every octave of <= 4096 pixels.
//
// Roofline: HBM/L2-bound in the limit. Algorithmic bytes per layer = 8*w*h (read once, write once).
#include <cuda.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "common.cuh"

namespace b200sift {

namespace {

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
    const float v11 = __ldg(in + (size_t)y1 * in_pitch + x1), v21 = __ldg(in + (size_t)y1 * in_pitch + x2);
    const float v12 = __ldg(in + (size_t)y2 * in_pitch + x1), v22 = __ldg(in + (size_t)y2 * in_pitch + x2);
    const float a = __fmaf_rn(v11, x2e, __fmul_rn(v21, xf));
    const float r12 = __fmul_rn(__fmaf_rn(v12, x2e, __fmul_rn(v22, xf)), yf);
    return __fmaf_rn(a, y2e, r12);
}

// The same resample, split: the per-axis part (indices and fraction depend on one coordinate only) ...
__device__ __forceinline__ void bilinear_axis(int i, int in_dim, int out_dim, int& i1, int& i2, float& f) {
    const float coord = __fsub_rn(__fdiv_rn(__fmul_rn(__fadd_rn((float)i, 0.5f), (float)in_dim), (float)out_dim), 0.5f);
    const int mx = in_dim - 1;
    i1 = min(mx, max(0, __float2int_rd(coord)));
    f = __fsub_rn(coord, truncf(coord));
    i2 = f == 0.f ? i1 + 1 : __float2int_ru(coord);
    i2 = min(mx, max(0, i2));
}
// ... and the combination of the four samples (identical operations in identical order)
__device__ __forceinline__ float bilinear_combine(float v11, float v21, float v12, float v22, float xf, float yf) {
    const float x2e = __fsub_rn(1.f, xf), y2e = __fsub_rn(1.f, yf);
    const float a = __fmaf_rn(v11, x2e, __fmul_rn(v21, xf));
    const float r12 = __fmul_rn(__fmaf_rn(v12, x2e, __fmul_rn(v22, xf)), yf);
    return __fmaf_rn(a, y2e, r12);
}

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---------------------------------------------------------------- fused blur chain
struct ChainParams {
    const float* src;        // plain: the layer to blur; resample: layer n of the previous octave
    int src_pitch, src_w, src_h;
    float* base_out;         // resample only: the octave's base layer
    float* out[3];           // layer written by stage 1..3
    int pitch, w, h;         // this octave
    Taps taps[3];
};

__host__ __device__ constexpr int round_up4(int v) { return (v + 3) & ~3; }

// Geometry shared by kernel and launcher. All shared-memory buffers use ONE coordinate frame: buffer column
// x - BX, buffer row y - BY with BX = (tx0 - E0 - 6) rounded down to a multiple of 4 (so that image columns that are
// multiples of 4 sit on 16-byte boundaries and every LDS.128 window of the horizontal pass stays inside the buffer)
// and BY = ty0 - E0, E0 = R1 + R2 + R3 the halo of stage 0.
template <int TW, int TH, int RB, int R1, int R2, int R3>
struct ChainGeom {
    static constexpr int K = (R2 < 0) ? 1 : (R3 < 0 ? 2 : 3);
    static constexpr int E0 = R1 + (R2 < 0 ? 0 : R2) + (R3 < 0 ? 0 : R3);
    static constexpr int E1 = E0 - R1;
    static constexpr int SW = round_up4(TW + 2 * E0 + 16);   // row stride of every buffer (floats)
    static constexpr int SH = TH + 2 * E0;                   // rows of the stage-0 region
    static constexpr int SHV = TH + 2 * E1;                  // rows of the largest row-filtered intermediate
    static constexpr int kLutRows = SH + RB + 8, kLutCols = SW + 8;
    static constexpr int kResampleTab = 3 * (SW + SH);       // per column / row of the stage-0 region: first index, second index, fraction
    // [S even | S odd (K > 1)] [V] [slack rows: partial register blocks read past their region] [row / column tables]
    static constexpr size_t kFloats = (size_t)SW * SH * (K > 1 ? 2 : 1) + (size_t)SW * SHV + (size_t)SW * RB;
    static constexpr size_t kSmem = kFloats * sizeof(float) + (size_t)(kLutRows + kLutCols + kResampleTab) * sizeof(int) + 16;
};

// Vertical pass of one stage: V(x, y) = chain over rows of S(x, reflect(y + d)) for the columns [xa, xb) and the rows
// [ya, yb) (image coordinates). A thread owns one column of one RB-row block: RB + 2R values in registers (compile
// time shared-memory offsets in the interior; through the reflect table at the top / bottom image border), RB
// in-order chains.
template <int R, int RB, int SW, int NT>
__device__ __forceinline__ void chain_vpass(const float* __restrict__ S, float* __restrict__ V, const Taps& t, int xa, int xb,
                                            int ya, int yb, int BX, int BY, int BYV, bool border, const int* __restrict__ rowlut, int tid) {
    constexpr int NTAP = 2 * R + 1;
    const int ncols = xb - xa;
    const int nblk = (yb - ya + RB - 1) / RB;
    const float inv_ncols = __frcp_rn((float)ncols);
    const int items = ncols * nblk;
    for (int it = tid; it < items; it += NT) {
        // it / ncols without an integer division: items < 2^16, (it + 0.5) / ncols is never within rounding distance of an integer
        const int blk = __float2int_rd(((float)it + 0.5f) * inv_ncols);
        const int c = xa + (it - blk * ncols) - BX;
        const int y0 = ya + blk * RB;
        float v[RB + 2 * R];
        if (!border) {
            const float* p = S + (y0 - R - BY) * SW + c;
#pragma unroll
            for (int i = 0; i < RB + 2 * R; i++) v[i] = p[i * SW];
        } else {
            const int* lut = rowlut + blk * RB;
#pragma unroll
            for (int i = 0; i < RB + 2 * R; i++) v[i] = S[lut[i] + c];
        }
        float* dst = V + (y0 - BYV) * SW + c;
#pragma unroll
        for (int r = 0; r < RB; r++) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < NTAP; k++) acc = __fmaf_rn(v[r + k], t.v[k], acc);
            if (y0 + r < yb) dst[r * SW] = acc;
        }
    }
}

// Horizontal pass: out(x, y) = chain over columns of V(reflect(x + d), y) for the quads [xq0, xq1) (multiples of 4) and
// rows [ya, yb). A thread produces 4 consecutive outputs from 4 + 2R row-filtered values (LDS.128 in the interior,
// through the reflect table at the left / right image border), stores them into the next stage's buffer (Snext, may
// be null) and -- for the part inside the CTA's tile and the image -- into the layer `out`.
template <int R, int SW, int NT>
__device__ __forceinline__ void chain_hpass(const float* __restrict__ V, float* __restrict__ Snext, float* __restrict__ out, int pitch,
                                            const Taps& t, int xq0, int xq1, int ya, int yb, int BX, int BY, int BYV, bool border,
                                            const int* __restrict__ collut, int tx0, int tx1, int ty0, int ty1, int tid) {
    constexpr int NTAP = 2 * R + 1;
    constexpr int RP = round_up4(R);
    const int nq = (xq1 - xq0) >> 2;
    const float inv_nq = __frcp_rn((float)nq);
    const int items = nq * (yb - ya);
    for (int it = tid; it < items; it += NT) {
        const int row = __float2int_rd(((float)it + 0.5f) * inv_nq);
        const int x = xq0 + 4 * (it - row * nq);
        const int y = ya + row;
        const float* vrow = V + (y - BYV) * SW;
        float o[4];
        if (!border) {
            constexpr int NV = (2 * RP + 4) / 4;
            float q[NV * 4];
            const float4* src = reinterpret_cast<const float4*>(vrow + (x - RP - BX));
#pragma unroll
            for (int m = 0; m < NV; m++) {
                const float4 f = src[m];
                q[4 * m] = f.x; q[4 * m + 1] = f.y; q[4 * m + 2] = f.z; q[4 * m + 3] = f.w;
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < NTAP; k++) acc = __fmaf_rn(q[RP - R + j + k], t.v[k], acc);
                o[j] = acc;
            }
        } else {
            float q[4 + 2 * R];
            const int* lut = collut + (x - xq0);
#pragma unroll
            for (int i = 0; i < 4 + 2 * R; i++) q[i] = vrow[lut[i]];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < NTAP; k++) acc = __fmaf_rn(q[j + k], t.v[k], acc);
                o[j] = acc;
            }
        }
        if (Snext) *reinterpret_cast<float4*>(Snext + (y - BY) * SW + (x - BX)) = make_float4(o[0], o[1], o[2], o[3]);
        if (y >= ty0 && y < ty1 && x >= tx0 && x < tx1) {   // tx0, tx1 and x are multiples of 4: a quad is inside or outside as a whole
            float* dst = out + (size_t)y * pitch + x;
            if (x + 3 < tx1) {
                *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
            } else {
                for (int j = 0; j < 4 && x + j < tx1; j++) dst[j] = o[j];
            }
        }
    }
}

template <bool DOWN, int NT, int TW, int TH, int RB, int R1, int R2, int R3>
__global__ void __launch_bounds__(NT, 2)
blur_chain_kernel(const __grid_constant__ ChainParams p, const __grid_constant__ CUtensorMap src_map) {
    using G = ChainGeom<TW, TH, RB, R1, R2, R3>;
    constexpr int K = G::K, E0 = G::E0, SW = G::SW, SH = G::SH;
    extern __shared__ __align__(128) float smem[];
    float* S0 = smem;
    float* S1 = smem + (size_t)SW * SH;                                   // only when K > 1
    float* Vb = smem + (size_t)SW * SH * (K > 1 ? 2 : 1);
    int* rowlut = reinterpret_cast<int*>(smem + G::kFloats);
    int* collut = rowlut + G::kLutRows;
    int* rtab = collut + G::kLutCols;                                     // resample tables (DOWN only)
    uint64_t* bar = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(rtab + G::kResampleTab) + 7) & ~uintptr_t(7));

    const int tid = threadIdx.x;
    const int w = p.w, h = p.h;
    const int tx0 = blockIdx.x * TW, ty0 = blockIdx.y * TH;
    const int tx1 = min(w, tx0 + TW), ty1 = min(h, ty0 + TH);  // tx1 may be ragged; tx0 is a multiple of 4
    const int BX = (tx0 - E0 - 6) & ~3, BY = ty0 - E0;
    const int BYV = ty0 - G::E1;   // row origin of the row-filtered intermediate (it never needs more than the stage-1 rows)

    // ---- stage 0: the region of the source this tile depends on -> S0
    {
        const int xa = max(0, tx0 - E0), xb = min(w, tx0 + TW + E0);
        const int ya = max(0, ty0 - E0), yb = min(h, ty0 + TH + E0);
        if constexpr (DOWN) {
            // per-column / per-row terms of the resample once per CTA, then four pixels (16 loads) in flight per thread
            const int ncols = xb - xa, nrows = yb - ya;
            int* cx1 = rtab; int* cx2 = rtab + SW; float* cxf = reinterpret_cast<float*>(rtab + 2 * SW);
            int* ry1 = rtab + 3 * SW; int* ry2 = ry1 + SH; float* ryf = reinterpret_cast<float*>(ry1 + 2 * SH);
            for (int j = tid; j < ncols + nrows; j += NT) {
                int i1, i2;
                float f;
                if (j < ncols) {
                    bilinear_axis(xa + j, p.src_w, w, i1, i2, f);
                    cx1[j] = i1; cx2[j] = i2; cxf[j] = f;
                } else {
                    bilinear_axis(ya + j - ncols, p.src_h, h, i1, i2, f);
                    ry1[j - ncols] = i1 * p.src_pitch; ry2[j - ncols] = i2 * p.src_pitch; ryf[j - ncols] = f;
                }
            }
            __syncthreads();
            const float inv = __frcp_rn((float)ncols);
            const int items = ncols * nrows;
            for (int it0 = tid; it0 < items; it0 += 4 * NT) {
                float v[4][4], xf[4], yf[4];
                int cc[4], rr[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int it = min(it0 + u * NT, items - 1);
                    rr[u] = __float2int_rd(((float)it + 0.5f) * inv);
                    cc[u] = it - rr[u] * ncols;
                    const float* r1 = p.src + ry1[rr[u]];
                    const float* r2 = p.src + ry2[rr[u]];
                    const int x1 = cx1[cc[u]], x2 = cx2[cc[u]];
                    v[u][0] = __ldg(r1 + x1); v[u][1] = __ldg(r1 + x2); v[u][2] = __ldg(r2 + x1); v[u][3] = __ldg(r2 + x2);
                    xf[u] = cxf[cc[u]]; yf[u] = ryf[rr[u]];
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (it0 + u * NT >= items) break;
                    const int x = xa + cc[u], y = ya + rr[u];
                    const float val = bilinear_combine(v[u][0], v[u][1], v[u][2], v[u][3], xf[u], yf[u]);
                    S0[(y - BY) * SW + (x - BX)] = val;
                    if (x >= tx0 && x < tx1 && y >= ty0 && y < ty1) p.base_out[(size_t)y * p.pitch + x] = val;
                }
            }
        } else {
            // ONE TMA tile load brings the whole SW x SH box of the source layer (origin (BX, BY), zero fill outside the
            // image) into S0: no address arithmetic, no registers, full memory-level parallelism
            if (tid == 0) {
                asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(bar)));
                asm volatile("fence.mbarrier_init.release.cluster;");
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"((uint32_t)(SW * SH * sizeof(float))) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(smem_addr(S0)), "l"(&src_map), "r"(BX), "r"(BY), "r"(smem_addr(bar)) : "memory");
            }
            __syncthreads();  // the barrier is initialised before anybody polls it
            uint32_t done = 0;
            for (uint32_t spins = 0; !done; spins++) {
                if (spins > (1u << 28)) __trap();  // a lost TMA must fail loudly, never hang the GPU
                asm volatile(
                    "{\n\t.reg .pred q;\n\t"
                    "mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\t"
                    "selp.u32 %0, 1, 0, q;\n\t}"
                    : "=r"(done) : "r"(smem_addr(bar)), "r"(0u) : "memory");
            }
            (void)xa; (void)xb; (void)ya; (void)yb;
        }
    }

    // ---- stages 1..K
    auto stage = [&](auto rtag, int k, const Taps& t, const float* S, float* Snext, float* out, int Ek) {
        constexpr int R = decltype(rtag)::value;
        // stage k is needed on the tile grown by Ek (the radii of the later stages), clipped to the image; its row-filtered
        // intermediate on the columns of that region grown by R (clipped): everything a reflected read can land on
        const int xa = max(0, tx0 - Ek), xb = min(w, tx0 + TW + Ek);
        const int ya = max(0, ty0 - Ek), yb = min(h, ty0 + TH + Ek);
        const int vxa = max(0, xa - R), vxb = min(w, xb + R);
        const bool vborder = (ya - R < 0) || (yb + R > h);
        const int xq0 = xa & ~3, xq1 = (xb + 3) & ~3;
        const bool hborder = (xq0 - R < 0) || (xq1 + R > w);
        __syncthreads();   // S complete (previous stage / stage 0); the tables of the previous stage are no longer read
        if (vborder) {
            const int nrows = ((yb - ya + RB - 1) / RB) * RB + 2 * R;
            for (int j = tid; j < nrows; j += NT) rowlut[j] = min(max(reflect101(ya - R + j, h) - BY, 0), SH - 1) * SW;
        }
        if (hborder) {
            const int ncols = (xq1 - xq0) + 2 * R;
            for (int j = tid; j < ncols; j += NT) collut[j] = min(max(reflect101(xq0 - R + j, w) - BX, 0), SW - 1);
        }
        if (vborder || hborder) __syncthreads();
        chain_vpass<R, RB, SW, NT>(S, Vb, t, vxa, vxb, ya, yb, BX, BY, BYV, vborder, rowlut, tid);
        __syncthreads();
        chain_hpass<R, SW, NT>(Vb, Snext, out, p.pitch, t, xq0, xq1, ya, yb, BX, BY, BYV, hborder, collut, tx0, tx1, ty0, ty1, tid);
        (void)k;
    };
    stage(std::integral_constant<int, R1>{}, 1, p.taps[0], S0, K > 1 ? S1 : nullptr, p.out[0], E0 - R1);
    if constexpr (K >= 2) stage(std::integral_constant<int, (R2 < 0 ? 1 : R2)>{}, 2, p.taps[1], S1, K > 2 ? S0 : nullptr, p.out[1], E0 - R1 - R2);
    if constexpr (K >= 3) stage(std::integral_constant<int, (R3 < 0 ? 1 : R3)>{}, 3, p.taps[2], S0, nullptr, p.out[2], 0);
}

// ---------------------------------------------------------------- single-stage blur (octave 0, unfused fallbacks)
// One Gaussian layer per launch: vertical pass straight from global memory into registers (16 + 2R rows per thread, all
// loads up front), the row-filtered tile in shared memory (<= 40 KB), horizontal pass from LDS.128 windows, STG.128.
// Same-box A/B on a B200 at 1920x1200 (profiles/r2_pyramid_ab.md) against two replacements written in round 2 -- a
// TMA-staged 128x64 tile kernel (~107 KB of shared memory per CTA) and a streaming column-strip kernel (register sliding
// window, 10 KB of shared memory, 14% occupancy, long-scoreboard bound): octave 0 takes 65 us with this kernel, 67 us with
// the tile kernel, 88 us with the streaming one, at the same 2.8-5.4 M warp instructions per launch; both were removed.
constexpr int kVhTW = 128;      // output columns per CTA (one warp row = 32 lanes x 4 outputs)
constexpr int kVhRB = 16;       // output rows per vertical register block
constexpr int kVhThreads = 320;  // 10 warps: one vertical-pass sweep covers the 128 + 2R columns of two row blocks

// One CTA: TH = nb*kVhRB output rows x kVhTW output columns.
//  pass 1 (vertical):   each thread owns one column of one row block, keeps kVhRB+2R input pixels
//                       in registers (all loads issued up front) and emits kVhRB row-filtered
//                       values into shared memory. Columns cover the tile plus an R-wide halo,
//                       taken at reflect-101 positions, because the horizontal pass of the
//                       reference reads the row-filtered image at reflected columns.
//  pass 2 (horizontal): each lane produces 4 consecutive outputs from 4+2R shared-memory values
//                       fetched as conflict-free LDS.128, stores one STG.128.
template <int R>
__global__ void __launch_bounds__(kVhThreads)
blur_vh_kernel(const float* __restrict__ in, int in_pitch, float* __restrict__ out, int out_pitch,
               int w, int h, const __grid_constant__ Taps taps, int nb) {
    constexpr int RP = (R + 3) & ~3;       // halo rounded up so that rows stay 16-byte aligned
    constexpr int TWP = kVhTW + 2 * RP;      // shared row length
    constexpr int NC = kVhTW + 2 * R;        // columns that are really needed
    constexpr int NT = 2 * R + 1;
    extern __shared__ __align__(16) float tmp[];  // [nb*kVhRB][TWP]

    const int x0 = blockIdx.x * kVhTW;
    const int y0 = blockIdx.y * nb * kVhRB;
    const int tid = threadIdx.x;

    for (int t = tid; t < NC * nb; t += kVhThreads) {
        const int b = t / NC;
        const int c = (RP - R) + (t - b * NC);
        const int x = reflect101(x0 - RP + c, w);
        const int ybase = y0 + b * kVhRB - R;
        float v[kVhRB + 2 * R];
        if (ybase >= 0 && ybase + kVhRB + 2 * R <= h) {
            // one 64-bit pointer stepped by the row pitch: a single IMAD.WIDE per load instead of the four address
            // instructions (IADD3 / IMAD.X / LEA / LEA.HI.X) an element-indexed p[i * pitch] compiles to
            const float* p = in + (size_t)ybase * in_pitch + x;
#pragma unroll
            for (int i = 0; i < kVhRB + 2 * R; i++) {
                v[i] = __ldg(p);
                asm volatile("mad.wide.s32 %0, %1, 4, %0;" : "+l"(p) : "r"(in_pitch));
            }
        } else {
#pragma unroll
            for (int i = 0; i < kVhRB + 2 * R; i++) v[i] = __ldg(in + (size_t)reflect101(ybase + i, h) * in_pitch + x);
        }
        float* dst = tmp + (size_t)(b * kVhRB) * TWP + c;
#pragma unroll
        for (int r = 0; r < kVhRB; r++) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < NT; k++) acc = __fmaf_rn(v[r + k], taps.v[k], acc);
            dst[r * TWP] = acc;
        }
    }
    __syncthreads();

    const int warp = tid >> 5, lane = tid & 31;
    const int TH = nb * kVhRB;
    const int xo = 4 * lane;
    const int x = x0 + xo;
    if (x >= w) return;
    for (int row = warp; row < TH; row += kVhThreads / 32) {
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
void launch_vh_r(const float* in, int in_pitch, float* out, int out_pitch, int w, int h, const Taps& taps, cudaStream_t s) {
    constexpr int RP = (R + 3) & ~3;
    constexpr int TWP = kVhTW + 2 * RP;
    // Taller tiles amortise the vertical halo on big images; short tiles spread small images
    // over more SMs.
    const int gx = (w + kVhTW - 1) / kVhTW;
    int nb = 4;
    while (nb > 1 && gx * ((h + nb * kVhRB - 1) / (nb * kVhRB)) < 2 * 148) nb >>= 1;
    dim3 grid(gx, (h + nb * kVhRB - 1) / (nb * kVhRB));
    size_t smem = (size_t)nb * kVhRB * TWP * sizeof(float);
    blur_vh_kernel<R><<<grid, kVhThreads, smem, s>>>(in, in_pitch, out, out_pitch, w, h, taps, nb);
}


template <int R>
struct VhTable {
    static void launch(int r, const float* in, int in_pitch, float* out, int out_pitch, int w, int h, const Taps& taps, cudaStream_t s) {
        if (r == R) launch_vh_r<R>(in, in_pitch, out, out_pitch, w, h, taps, s);
        else VhTable<R - 1>::launch(r, in, in_pitch, out, out_pitch, w, h, taps, s);
    }
};
template <>
struct VhTable<0> {
    static void launch(int, const float*, int, float*, int, int, int, const Taps&, cudaStream_t) {}
};
constexpr int kMaxVhR = 16;

// ---------------------------------------------------------------- generic single-stage blur (any radius <= 42)
constexpr int kGenTW = 64, kGenTH = 32, kGenThreads = 256;
constexpr int kGenMaxR = (kMaxTaps - 1) / 2;

__global__ void __launch_bounds__(kGenThreads)
blur_generic_kernel(const float* __restrict__ in, int in_pitch, float* __restrict__ out, int out_pitch, int w, int h,
                    const __grid_constant__ Taps taps) {
    extern __shared__ __align__(128) float smem[];
    const int R = taps.n / 2;
    const int CW = kGenTW + 2 * R;           // columns of the row-filtered tile
    float* V = smem;                         // [kGenTH][CW]
    const int tid = threadIdx.x;
    const int tx0 = blockIdx.x * kGenTW, ty0 = blockIdx.y * kGenTH;
    for (int it = tid; it < CW * kGenTH; it += kGenThreads) {
        const int r = it / CW, c = it - r * CW;
        const int y = ty0 + r;
        const int x = reflect101(tx0 - R + c, w);   // the horizontal pass reads the row-filtered image at reflected columns
        float acc = 0.f;
        if (y < h)
            for (int k = 0; k < taps.n; k++) acc = __fmaf_rn(__ldg(in + (size_t)reflect101(y - R + k, h) * in_pitch + x), taps.v[k], acc);
        V[it] = acc;
    }
    __syncthreads();
    for (int it = tid; it < kGenTW * kGenTH; it += kGenThreads) {
        const int r = it / kGenTW, c = it - r * kGenTW;
        const int x = tx0 + c, y = ty0 + r;
        if (x >= w || y >= h) continue;
        const float* row = V + r * CW + c;
        float acc = 0.f;
        for (int k = 0; k < taps.n; k++) acc = __fmaf_rn(row[k], taps.v[k], acc);
        out[(size_t)y * out_pitch + x] = acc;
    }
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
// Octaves of at most kTailPixels pixels are pure launch latency as separate kernels. ONE CTA walks all of them
// instead: the current layer and the row-filtered intermediate live in shared memory, every layer is also written to
// its place in the Gaussian stack. Arithmetic is the same in-order fmaf chain as above.
struct TailParams {
    Taps taps[kMaxLayers];
    Taps init_taps;
    const float* input;      // the frame (only read when the tail starts at octave 0)
    int in_pitch;
    float* gauss[kMaxOctaves];
    int w[kMaxOctaves], h[kMaxOctaves], pitch[kMaxOctaves];
    int first, nOct, nScales;
    int l_begin, l_end;
};

constexpr int kTailThreads = 1024;
constexpr int kTailR = kGenMaxR;  // largest blur radius the reflect tables cover

// One blur layer of the tail: vertical pass s_cur -> s_tmp, horizontal pass s_tmp -> s_cur (+ global).
// R > 0: compile-time radius, both passes fully unrolled; R == 0: the radius is read from the taps.
template <int R>
__device__ __forceinline__ void tail_layer(const Taps& t, float* s_cur, float* s_tmp, const int* s_rtab, const int* s_ctab,
                                           float* __restrict__ dst, int w, int pitch, int n, int tid) {
    // i / w without an integer division: n <= 4096, so the float floor of (i + 0.5) / w is exact
    const float inv_w = __frcp_rn((float)w);
    if constexpr (R > 0) {
        constexpr int NT = 2 * R + 1;
        float tap[NT];
#pragma unroll
        for (int k = 0; k < NT; k++) tap[k] = t.v[k];
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
    } else {
        const int nt = t.n, r = t.n / 2;
        for (int i = tid; i < n; i += kTailThreads) {
            const int y = __float2int_rd(((float)i + 0.5f) * inv_w), x = i - y * w;
            const int* rt = s_rtab + (y - r + kTailR);
            float acc = 0.f;
            for (int k = 0; k < nt; k++) acc = __fmaf_rn(s_cur[rt[k] + x], t.v[k], acc);
            s_tmp[i] = acc;
        }
        __syncthreads();
        for (int i = tid; i < n; i += kTailThreads) {
            const int y = __float2int_rd(((float)i + 0.5f) * inv_w), x = i - y * w;
            const int* ct = s_ctab + (x - r + kTailR);
            const float* row = s_tmp + y * w;
            float acc = 0.f;
            for (int k = 0; k < nt; k++) acc = __fmaf_rn(row[ct[k]], t.v[k], acc);
            s_cur[i] = acc;
            dst[(size_t)y * pitch + x] = acc;
        }
    }
    __syncthreads();  // also orders this CTA's global writes of layer n before the next octave reads them
}

__device__ __forceinline__ void tail_layer_any(const Taps& t, float* s_cur, float* s_tmp, const int* s_rtab, const int* s_ctab,
                                               float* __restrict__ dst, int w, int pitch, int n, int tid) {
    switch (t.n / 2) {  // the radii of the default configuration get unrolled code
        case 4: tail_layer<4>(t, s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
        case 5: tail_layer<5>(t, s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
        case 6: tail_layer<6>(t, s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
        case 8: tail_layer<8>(t, s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
        case 10: tail_layer<10>(t, s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
        default: tail_layer<0>(t, s_cur, s_tmp, s_rtab, s_ctab, dst, w, pitch, n, tid); break;
    }
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
    // l_begin == 0: the octaves depend on each other (one CTA walks them all); l_begin > 0: every octave continues from
    // its own stored layer l_begin - 1, CTA b takes octave first + b
    const int o_lo = tp.l_begin > 0 ? tp.first + (int)blockIdx.x : tp.first;
    const int o_hi = tp.l_begin > 0 ? o_lo + 1 : tp.nOct;
    for (int o = o_lo; o < o_hi; o++) {
        const int w = tp.w[o], h = tp.h[o], pitch = tp.pitch[o], n = w * h;
        float* g = tp.gauss[o];
        const size_t ls = (size_t)pitch * h;
        for (int j = tid; j < h + 2 * kTailR; j += kTailThreads) s_rtab[j] = reflect101(j - kTailR, h) * w;
        for (int j = tid; j < w + 2 * kTailR; j += kTailThreads) s_ctab[j] = reflect101(j - kTailR, w);
        if (tp.l_begin > 0) {    // continue from a layer that an earlier launch left in the stack
            const float* src = g + ls * (tp.l_begin - 1);
            for (int i = tid; i < n; i += kTailThreads) s_cur[i] = src[(size_t)(i / w) * pitch + i % w];
            __syncthreads();
        } else if (o == 0) {     // the frame itself: initial blur (interface/Detector.cu:253-258) -> layer 0
            for (int i = tid; i < n; i += kTailThreads) s_cur[i] = tp.input[(size_t)(i / w) * tp.in_pitch + i % w];
            __syncthreads();
            tail_layer_any(tp.init_taps, s_cur, s_tmp, s_rtab, s_ctab, g, w, pitch, n, tid);
        } else {                 // bilinear resample of the previous octave's layer n (written by this CTA or by the launch before)
            const float* src = tp.gauss[o - 1] + (size_t)tp.pitch[o - 1] * tp.h[o - 1] * tp.nScales;
            for (int i = tid; i < n; i += kTailThreads) {
                const int y = i / w, x = i - y * w;
                const float v = bilinear_sample(src, tp.pitch[o - 1], tp.w[o - 1], tp.h[o - 1], w, h, x, y);
                s_cur[i] = v;
                g[(size_t)y * pitch + x] = v;
            }
            __syncthreads();
        }
        for (int l = max(1, tp.l_begin); l < tp.l_end; l++)
            tail_layer_any(tp.taps[l], s_cur, s_tmp, s_rtab, s_ctab, g + ls * l, w, pitch, n, tid);
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

// ---- launch tables -------------------------------------------------------------------------------------------
extern bool g_pyramid_error;
// Fused chains: 256 threads, 8-row register blocks, 64 x 32 tiles. Smaller tiles for the small octaves (32 x 16 at
// 480 x 300, 16 x 8 below: a CTA's latency grows with its tile plus the 15 / 18 pixel halo, and a 64 x 32 grid leaves most
// SMs idle there) were measured on a B200 at 1920x1200: the three launches got 4-5.5 us shorter each under ncu, the
// pyramid 4 us shorter inside the frame graph (0.151 -> 0.147 ms), and the pipelined throughput fell from 3700 to 3456
// frames/s (three times as many CTAs queue behind the other frames' descriptor CTAs, twice the halo work). The smaller
// shapes stay compiled behind B200SIFT_CHAIN_SHAPES=1 so that the trade can be re-measured (profiles/r2_pyramid_ab.md).
constexpr int kFuseThreads = 256, kFuseRB = 8;
constexpr int kNumShapes = 3;
template <int I> struct Shape;
template <> struct Shape<0> { static constexpr int TW = 64, TH = 32; };
template <> struct Shape<1> { static constexpr int TW = 32, TH = 16; };
template <> struct Shape<2> { static constexpr int TW = 16, TH = 8; };

int g_small_shapes = -1;  // development switch B200SIFT_CHAIN_SHAPES=1: the largest shape that still gives 148 CTAs

int pick_shape(int w, int h) {
    if (g_small_shapes < 0) {
        const char* e = std::getenv("B200SIFT_CHAIN_SHAPES");
        g_small_shapes = (e && e[0] == '1') ? 1 : 0;
    }
    if (!g_small_shapes) return 0;
    const int tw[kNumShapes] = {Shape<0>::TW, Shape<1>::TW, Shape<2>::TW}, th[kNumShapes] = {Shape<0>::TH, Shape<1>::TH, Shape<2>::TH};
    for (int i = 0; i < kNumShapes; i++)
        if (((w + tw[i] - 1) / tw[i]) * ((h + th[i] - 1) / th[i]) >= 148) return i;
    return kNumShapes - 1;
}

// Tensor map of a w x h float layer with row pitch `pitch` floats; box = the CTA's stage-0 region. The encode function
// comes from the driver through the runtime's entry-point query (no link-time dependency on libcuda).
bool encode_layer_map(CUtensorMap* map, const float* base, int w, int h, int pitch, int box_w, int box_h) {
    using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) return false;
        encode = reinterpret_cast<EncodeFn>(fn);
    }
    const cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h};
    const cuuint64_t strides[1] = {(cuuint64_t)pitch * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)box_w, (cuuint32_t)box_h};
    const cuuint32_t estr[2] = {1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <bool DOWN, int NT, int TW, int TH, int RB, int R1, int R2, int R3>
void launch_chain_inst(const ChainParams& p, cudaStream_t s) {
    using G = ChainGeom<TW, TH, RB, R1, R2, R3>;
    static_assert(G::SW <= 256 && G::SH <= 256, "TMA box dimensions");
    dim3 grid((p.w + TW - 1) / TW, (p.h + TH - 1) / TH);
    CUtensorMap map;
    std::memset(&map, 0, sizeof(map));
    if (!DOWN && !encode_layer_map(&map, p.src, p.w, p.h, p.src_pitch, G::SW, G::SH)) {
        g_pyramid_error = true;  // reported by pyramid_failed() (engine.cu turns it into an error of the enqueueing call)
        return;
    }
    blur_chain_kernel<DOWN, NT, TW, TH, RB, R1, R2, R3><<<grid, NT, G::kSmem, s>>>(p, map);
}
template <bool DOWN, int NT, int TW, int TH, int RB, int R1, int R2, int R3>
cudaError_t setup_chain_inst() {
    using G = ChainGeom<TW, TH, RB, R1, R2, R3>;
    return cudaFuncSetAttribute(blur_chain_kernel<DOWN, NT, TW, TH, RB, R1, R2, R3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::kSmem);
}

bool g_pyramid_error = false;

// the fused chains of the default configuration (sigma 1.6, numOctaveLayers 3): resample + 9/11/13 taps (layers 1..3 of
// an octave, from layer 3 of the octave before) and 17/21 taps (layers 4, 5), each in every tile shape
template <int I>
struct ChainTable {
    using S = Shape<I>;
    static void launch(int shape, bool down, const ChainParams& p, cudaStream_t s) {
        if (shape != I) return ChainTable<I + 1>::launch(shape, down, p, s);
        if (down) launch_chain_inst<true, kFuseThreads, S::TW, S::TH, kFuseRB, 4, 5, 6>(p, s);
        else launch_chain_inst<false, kFuseThreads, S::TW, S::TH, kFuseRB, 8, 10, -1>(p, s);
    }
    static cudaError_t setup() {
        cudaError_t e = setup_chain_inst<true, kFuseThreads, S::TW, S::TH, kFuseRB, 4, 5, 6>();
        if (e == cudaSuccess) e = setup_chain_inst<false, kFuseThreads, S::TW, S::TH, kFuseRB, 8, 10, -1>();
        return e != cudaSuccess ? e : ChainTable<I + 1>::setup();
    }
};
template <>
struct ChainTable<kNumShapes> {
    static void launch(int, bool, const ChainParams&, cudaStream_t) {}
    static cudaError_t setup() { return cudaSuccess; }
};

bool is_radii(const BlurStage* st, int n, int a, int b, int c) {
    const int want[3] = {a, b, c};
    int k = 0;
    while (k < 3 && want[k] >= 0) k++;
    if (n != k) return false;
    for (int i = 0; i < n; i++)
        if (st[i].taps->n != 2 * want[i] + 1) return false;
    return true;
}

}  // namespace

bool pyramid_failed() {
    const bool f = g_pyramid_error;
    g_pyramid_error = false;
    return f;
}

cudaError_t pyramid_setup() {
    cudaError_t e = ChainTable<0>::setup();
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(blur_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)((kGenTW + 2 * kGenMaxR) * kGenTH * sizeof(float)));
}

int g_fuse_mode = -1;  // 1 (default): fused resample + blur chains for octaves >= 1; 0: one launch per layer

bool blur_chain_supported(bool from_prev_octave, const BlurStage* st, int n) {
    if (g_fuse_mode < 0) {  // development switch for same-box A/B runs: B200SIFT_FUSE=0 -> one launch per layer everywhere
        const char* e = std::getenv("B200SIFT_FUSE");
        g_fuse_mode = (e && e[0] == '0') ? 0 : 1;
    }
    if (!g_fuse_mode) return false;
    return from_prev_octave ? is_radii(st, n, 4, 5, 6) : is_radii(st, n, 8, 10, -1);
}

void launch_blur_chain(const float* src, int src_pitch, int src_w, int src_h, bool from_prev_octave, float* base_out,
                       const BlurStage* st, int n, int pitch, int w, int h, cudaStream_t s) {
    ChainParams p;
    p.src = src; p.src_pitch = src_pitch; p.src_w = src_w; p.src_h = src_h;
    p.base_out = base_out;
    p.pitch = pitch; p.w = w; p.h = h;
    for (int i = 0; i < 3; i++) {
        p.out[i] = i < n ? st[i].out : nullptr;
        if (i < n) p.taps[i] = *st[i].taps;
        else p.taps[i].n = 0;
    }
    ChainTable<0>::launch(pick_shape(w, h), from_prev_octave, p, s);
}

void launch_blur(const float* in, int in_pitch, float* out, int out_pitch, int w, int h, const Taps& taps, cudaStream_t s) {
    const int r = taps.n / 2;
    if ((taps.n & 1) && r >= 1 && r <= kMaxVhR) {
        VhTable<kMaxVhR>::launch(r, in, in_pitch, out, out_pitch, w, h, taps, s);
        return;
    }
    dim3 grid((w + kGenTW - 1) / kGenTW, (h + kGenTH - 1) / kGenTH);
    blur_generic_kernel<<<grid, kGenThreads, (size_t)(kGenTW + 2 * r) * kGenTH * sizeof(float), s>>>(in, in_pitch, out, out_pitch, w, h, taps);
}

void launch_downsample(const float* in, int in_pitch, int iw, int ih, float* out, int out_pitch, int ow, int oh, cudaStream_t s) {
    dim3 grid((ow + 31) / 32, (oh + 7) / 8);
    downsample_kernel<<<grid, 256, 0, s>>>(in, in_pitch, iw, ih, out, out_pitch, ow, oh);
}

void launch_pyramid_tail(const FrameParams& fp, const Taps* taps, const Taps& init_taps, const float* input, int in_pitch,
                         int first, int l_begin, int l_end, cudaStream_t s) {
    TailParams tp;
    for (int l = 0; l < kMaxLayers; l++) {
        if (l < fp.nLayers) tp.taps[l] = taps[l];
        else tp.taps[l].n = 0;
    }
    tp.init_taps = init_taps;
    tp.input = input;
    tp.in_pitch = in_pitch;
    for (int o = 0; o < kMaxOctaves; o++) {
        tp.gauss[o] = o < fp.nOct ? fp.oct[o].gauss : nullptr;
        tp.w[o] = o < fp.nOct ? fp.oct[o].w : 0;
        tp.h[o] = o < fp.nOct ? fp.oct[o].h : 0;
        tp.pitch[o] = o < fp.nOct ? fp.oct[o].pitch : 0;
    }
    tp.first = first;
    tp.nOct = fp.nOct;
    tp.nScales = fp.nScales;
    tp.l_begin = l_begin;
    tp.l_end = l_end;
    pyramid_tail_kernel<<<l_begin > 0 ? fp.nOct - first : 1, kTailThreads, 0, s>>>(tp);
}

void launch_u8_to_float(const uint8_t* in, size_t in_pitch, float* out, int out_pitch, int w, int h, cudaStream_t s) {
    dim3 grid((w + 1023) / 1024, h);
    u8_to_float_kernel<<<grid, 256, 0, s>>>(in, in_pitch, out, out_pitch, w, h);
}

}  // namespace b200sift
