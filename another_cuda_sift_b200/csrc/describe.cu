// Orientation assignment, keypoint expansion / compaction and the 128-D descriptor.
//
// Replaces sift_func/SiftOps.cu:237-376 (calOriHistMultiThread), :210-235 (collectKpts, second
// use), :389-623 (genDescriptorMultiThread and helpers), utils/CudaMemcpyUtils.cu (copyValidNum,
// collectKptsAndDescriptor) and the host round trip of interface/Detector.cu:536-604: all counts
// stay on the device and descriptors are written straight into the final output arrays.
//
// Per-sample arithmetic follows the reference's expression trees with the FMA contraction its
// sm_100a SASS shows. The one deliberate difference: the reference accumulates histogram votes
// with float shared-memory atomics, whose order -- and therefore rounding -- changes from run to
// run; here votes are accumulated in fixed point with native integer shared-memory atomics, which is
// order independent (deterministic) and lies inside the reference's own run-to-run envelope:
// orientation bins as 2^-32 in 64 bits (peak selection is a discrete decision, keep it as exact as
// possible), descriptor bins as 2^-15 in 32 bits (8 votes per sample: one ATOMS.ADD each).
#include <cstdlib>

#include "common.cuh"

namespace b200sift {

namespace {

constexpr float kFix = 4294967296.f;          // 2^32
constexpr double kUnfix = 1.0 / 4294967296.0;

// 64-bit fixed-point accumulate built from NATIVE 32-bit shared-memory atomics (ATOMS.ADD): add the low
// word, derive the carry from the returned old value, add high word + carry. The final 64-bit value is
// the exact sum mod 2^64 whatever the order, i.e. deterministic; a 64-bit or float shared atomicAdd would
// compile to a compare-and-swap loop that serialises the many lanes voting into the same bin.
__device__ __forceinline__ void fix_add(unsigned long long* slot, float v) {
    const long long q = __float2ll_rn(__fmul_rn(v, kFix));
    const unsigned lo = (unsigned)q;
    int hi = (int)(q >> 32);
    unsigned* p = reinterpret_cast<unsigned*>(slot);
    const unsigned old = atomicAdd(p, lo);
    hi += (old + lo < old) ? 1 : 0;
    if (hi) atomicAdd(p + 1, (unsigned)hi);
}
__device__ __forceinline__ float fix_get(unsigned long long v) { return (float)((double)(long long)v * kUnfix); }

// Descriptor bins: int32, unit 2^-15. Two's-complement wrap-around makes partial sums harmless; a final bin
// must lie within +-65536, far above what 0..255-valued pixels blurred with sigma >= 1.6 can produce
// (|gradient| < 170, weight <= 1, a few hundred samples per bin). Quantisation: <= 1.5e-5 per partial sum, i.e.
// ~1e-6 relative on a bin -- below the fp32 rounding noise of the reference's own float accumulation.
constexpr float kUnfix32 = 1.f / 32768.f;

// ---------------------------------------------------------------- orientation
// One 64-thread CTA per refined keypoint (flattened over the octaves, grid-stride). The patch loop is
// latency-bound (8 scattered loads + atan2f per sample), so every thread keeps two samples in flight.
constexpr int kOriThreads = 64;

struct OriSample {
    float val;   // weight * magnitude
    int bin;
};

__global__ void __launch_bounds__(kOriThreads)
orientation_kernel(const __grid_constant__ FrameParams fp) {
    __shared__ unsigned long long s_hist[kOriBins];
    __shared__ float s_raw[kOriBins];
    __shared__ float s_smooth[kOriBins];
    __shared__ int s_peak[kOriBins];
    const int tid = threadIdx.x;
    int total = 0;
    for (int i = fp.o_begin; i < fp.o_end; i++) total += fp.oct[i].counts[1];

    for (int g = blockIdx.x; g < total; g += gridDim.x) {
        int oi = fp.o_begin, k = g;
        while (k >= fp.oct[oi].counts[1]) { k -= fp.oct[oi].counts[1]; oi++; }
        const OctParams& o = fp.oct[oi];
        const int M = o.counts[1];
        const size_t ls = (size_t)o.pitch * o.h;
        const int h_dim = o.pitch;
        const float3 kp = o.coll[k];
        const float4 ft = o.collF[k];
        if (tid < kOriBins) s_hist[tid] = 0ull;
        __syncthreads();
        // SiftOps.cu:271-297
        // opt_upscale: SiftOps.cu:260,269 (mod = upscaled ? 0 : 1) with the halved size, and positions back in octave pixels
        const float scale = __fdiv_rn(__fmul_rn(1.5f, ft.y), (float)(1 << (oi + (fp.opt_upscale ? 0 : 1))));
        const float radius = roundf(__fmul_rn(3.f, scale));
        const float pos_mul = fp.opt_upscale ? 2.f : 1.f;
        const int base_px = (int)roundf(__fdiv_rn(__fmul_rn(kp.x, pos_mul), (float)(1 << oi)));
        const int base_py = (int)roundf(__fdiv_rn(__fmul_rn(kp.y, pos_mul), (float)(1 << oi)));
        int side = (int)__fmaf_rn(2.f, radius, 1.f);
        if (side > 1024) side = 1024;  // only a corrupt stale slot could ask for more
        const int total_dim = side * side;
        const float weight_factor = __fdiv_rn(-0.5f, __fmul_rn(scale, scale));
        int layer = (int)kp.z;
        layer = min(max(layer, 0), fp.nScales + 1);
        const float* g0 = o.gauss + ls * layer;  // DoG[layer] = G[layer+1] - G[layer]
        const float fh1 = (float)o.h - 1.f, fw1 = (float)o.w - 1.f;

        // idx / side without an integer division: (idx + 0.5) / side is at least 0.5 / side >= 4.9e-4 away from an integer,
        // the float product is off by < 1.3e-4 (idx < 2^20, side <= 1024), so the floor is exact
        const float inv_side = __frcp_rn((float)side);
        auto geometry = [&](int idx, int& px_off, int& py_off, const float*& p) -> bool {
            if (idx >= total_dim) return false;
            const int row = __float2int_rd(__fmul_rn((float)idx + 0.5f, inv_side));
            py_off = (int)__fsub_rn((float)row, radius);
            const int py = base_py + py_off;
            px_off = (int)__fsub_rn((float)(idx - row * side), radius);
            const int px = base_px + px_off;
            if (py < 1 || (float)py >= fh1 || px < 1 || (float)px >= fw1) return false;
            p = g0 + (size_t)py * h_dim + px;
            return true;
        };
        auto vote = [&](const float v[8], int px_off, int py_off) -> OriSample {
            const float dx = __fsub_rn(__fsub_rn(v[0], v[1]), __fsub_rn(v[2], v[3]));
            const float dy = __fsub_rn(__fsub_rn(v[4], v[5]), __fsub_rn(v[6], v[7]));
            const float mag = __fsqrt_rd(__fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
            const float weight = __expf(__fmul_rn(weight_factor, (float)(px_off * px_off + py_off * py_off)));
            // The angle only selects the bin. A polynomial atan2 (error < 2e-6 rad = 1.2e-5 bins, checked against
            // double precision over the gradient range) decides it whenever the bin coordinate is further than 1e-3
            // from a bin edge -- 99.8% of the samples; the others (and zero / non-finite gradients, for which the
            // comparisons below are false) take the reference's own expression. Same bins, a third of the instructions.
            int bin;
            {
                const float ax = fabsf(dx), ay = fabsf(dy);
                const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
                float rinv;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rinv) : "f"(mx));
                const float x = mn * rinv, x2 = x * x;
                float r = fmaf(fmaf(fmaf(fmaf(fmaf(-0.01172120f, x2, 0.05265332f), x2, -0.11643287f), x2, 0.19354346f), x2, -0.33262347f), x2, 0.99997726f) * x;
                r = ay > ax ? 1.57079632679f - r : r;
                r = dx < 0.f ? 3.14159265359f - r : r;
                r = dy < 0.f ? -r : r;
                float c = r * (18.f / 3.14159265359f);   // bins of 10 degrees
                c = c < 0.f ? c + 36.f : c;
                const float cf = floorf(c), f = c - cf;
                bin = (int)cf;
                if (!(f > 1e-3f && f < 1.f - 1e-3f && mx > 0.f)) {
                    float ori = __fmul_rn(__fdiv_rn(atan2f(dy, dx), 3.14159265358979323846f), 180.f);
                    if (ori < 0.f) ori = __fadd_rn(ori, 360.f);
                    bin = __float2int_rd(__fmul_rn(__fdiv_rn(ori, 360.f), 36.f)) % kOriBins;
                    if (bin < 0) bin += kOriBins;  // NaN guard; never taken for finite input
                }
            }
            OriSample r;
            r.val = __fmul_rn(weight, mag);
            r.bin = bin;
            return r;
        };
        auto load8 = [&](const float* p, float v[8]) {
            v[0] = __ldg(p + ls + 1); v[1] = __ldg(p + 1); v[2] = __ldg(p + ls - 1); v[3] = __ldg(p - 1);
            v[4] = __ldg(p + ls - h_dim); v[5] = __ldg(p - h_dim); v[6] = __ldg(p + ls + h_dim); v[7] = __ldg(p + h_dim);
        };

        for (int idx = tid; idx < total_dim; idx += 2 * kOriThreads) {
            int ax, ay, bx, by;
            const float *pa = g0, *pb = g0;
            const bool va = geometry(idx, ax, ay, pa);
            const bool vb = geometry(idx + kOriThreads, bx, by, pb);
            float a8[8], b8[8];
            if (va) load8(pa, a8);
            if (vb) load8(pb, b8);
            if (va) { const OriSample r = vote(a8, ax, ay); fix_add(s_hist + r.bin, r.val); }
            if (vb) { const OriSample r = vote(b8, bx, by); fix_add(s_hist + r.bin, r.val); }
        }
        __syncthreads();
        if (tid < kOriBins) s_raw[tid] = fix_get(s_hist[tid]);
        __syncthreads();
        // [1 4 6 4 1]/16 smoothing with the reference's association (SiftOps.cu:335)
        if (tid < kOriBins) {
            const int i = tid;
            const int l = i > 0 ? i - 1 : kOriBins - 1, r = i < kOriBins - 1 ? i + 1 : 0;
            const int r2 = r < kOriBins - 1 ? r + 1 : 0, l2 = l > 0 ? l - 1 : kOriBins - 1;
            const float a = __fmaf_rn(s_raw[i], 6.f, __fmul_rn(__fadd_rn(s_raw[l], s_raw[r]), 4.f));
            s_smooth[i] = __fmul_rn(__fadd_rn(__fadd_rn(a, s_raw[l2]), s_raw[r2]), 0.0625f);
        }
        __syncthreads();
        if (tid < kOriBins) {
            float mx = 0.f;
            for (int i = 0; i < kOriBins; i++) mx = fmaxf(mx, s_smooth[i]);
            const float thr = __fmul_rn(mx, 0.8f);
            const int i = tid;
            const int l = i > 0 ? i - 1 : kOriBins - 1, r = i < kOriBins - 1 ? i + 1 : 0;
            s_peak[i] = (s_smooth[i] > s_smooth[l] && s_smooth[i] > s_smooth[r] && s_smooth[i] >= thr) ? 1 : 0;
        }
        __syncthreads();
        // copy c of keypoint k goes to slot M + 36k + c, dropped once slot >= cap (quirk Q6)
        if (tid < kOriBins) {
            const int slot0 = k * kOriBins + M;
            int c = 0;
            for (int i = 0; i < tid; i++) c += s_peak[i];
            const int slot = slot0 + c;
            if (s_peak[tid] && slot < o.cap) {
                o.coll[slot] = kp;
                float4 f = ft;
                f.w = __fsub_rn(360.f, __fmul_rn(__fdiv_rn((float)tid, 36.f), 360.f));
                if (fp.opt_ori_interp) {
                    // opt-in (SiftOps.cu:359 "TODO: interpolate the peak"): parabola through the peak and its neighbours of the
                    // smoothed histogram, as OpenCV does (features2d/src/sift.simd.hpp:555-557)
                    const int l = tid > 0 ? tid - 1 : kOriBins - 1, r = tid < kOriBins - 1 ? tid + 1 : 0;
                    const float hl = s_smooth[l], hc = s_smooth[tid], hr = s_smooth[r];
                    float bin = (float)tid + 0.5f * (hl - hr) / (hl - 2.f * hc + hr);
                    bin = bin < 0.f ? (float)kOriBins + bin : (bin >= (float)kOriBins ? bin - (float)kOriBins : bin);
                    f.w = 360.f - (360.f / (float)kOriBins) * bin;
                    if (fabsf(f.w - 360.f) < 1.1920929e-7f) f.w = 0.f;
                }
                o.collF[slot] = f;
            }
            if (tid == kOriBins - 1) {
                const int emitted = c + s_peak[tid];
                o.oricnt[k] = max(0, min(emitted, o.cap - slot0));
            }
        }
        __syncthreads();
    }
}

// CTA (blockIdx.x, octave): prefix over the per-keypoint peak counts of its kScanBlock-keypoint segment ->
// cand/candF in (k, bin) order; the entry whose prefix equals the total T is not written (quirk Q5 again).
__global__ void __launch_bounds__(kScanBlock)
compact_oriented_kernel(const __grid_constant__ FrameParams fp) {
    const OctParams& o = fp.oct[fp.o_begin + blockIdx.y];
    const int M = o.counts[1];
    if (blockIdx.x > 0 && (int)blockIdx.x * kScanBlock >= M) return;
    const int total = segment_scan(
        M, blockIdx.x, [&](int k) { return o.oricnt[k]; },
        [&](int k, int v, int excl, int tot) {
            for (int c = 0; c < v; c++) {
                const int p = excl + c;
                float3 kp;
                float4 ft;
                if (p + 1 < tot) {
                    const int slot = M + k * kOriBins + c;
                    kp = o.coll[slot];
                    ft = o.collF[slot];
                    o.cand[p] = kp;
                    o.candF[p] = ft;
                } else {  // quirk Q5: entry T-1 keeps whatever an earlier frame (or stage) left there
                    kp = o.cand[p];
                    ft = o.candF[p];
                }
                // the descriptor stage reads this per-slot copy: cand/candF belong to the next frame from here on
                if (p < fp.numFeatures) { o.fin_kp[p] = kp; o.fin_ft[p] = ft; }
            }
        });
    if (blockIdx.x == 0 && threadIdx.x == 0) o.counts[2] = total;  // T
}

// ---------------------------------------------------------------- duplicate removal (opt-in)
// interface/Detector.cu:220 "TODO: remove duplicate". A keypoint is a duplicate when an EARLIER keypoint of the same octave
// has the same x, y, size and angle (what OpenCV's KeyPointsFilter::removeDuplicated compares); the first one is kept, the
// order of the survivors is unchanged. Two launches after compact_oriented: flags (each CTA compares its 256 entries with
// all earlier ones, tiled through shared memory), then a stable compaction into the slot's final list.
__global__ void __launch_bounds__(kScanBlock)
dedup_flag_kernel(const __grid_constant__ FrameParams fp) {
    const OctParams& o = fp.oct[fp.o_begin + blockIdx.y];
    const int T = min(o.counts[2], o.cap);
    const int i0 = blockIdx.x * kScanBlock;
    if (i0 >= T) return;
    __shared__ float4 s_key[kScanBlock];
    const int i = i0 + threadIdx.x;
    float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < T) {
        const float3 k = o.cand[i];
        const float4 f = o.candF[i];
        mine = make_float4(k.x, k.y, f.y, f.w);
    }
    bool dup = false;
    for (int j0 = 0; j0 <= i0; j0 += kScanBlock) {
        const int j = j0 + threadIdx.x;
        if (j < T) {
            const float3 k = o.cand[j];
            const float4 f = o.candF[j];
            s_key[threadIdx.x] = make_float4(k.x, k.y, f.y, f.w);
        }
        __syncthreads();
        const int n = min(kScanBlock, min(T, i) - j0);   // only entries before i
        for (int t = 0; t < n; t++) {
            const float4 q = s_key[t];
            dup |= (q.x == mine.x) & (q.y == mine.y) & (q.z == mine.z) & (q.w == mine.w);
        }
        __syncthreads();
    }
    if (i < T) o.flag[i] = dup ? 0 : 1;
}

__global__ void __launch_bounds__(kScanBlock)
dedup_compact_kernel(const __grid_constant__ FrameParams fp) {
    const OctParams& o = fp.oct[fp.o_begin + blockIdx.y];
    const int T = min(o.counts[2], o.cap);
    if (blockIdx.x > 0 && (int)blockIdx.x * kScanBlock >= T) return;
    const int total = segment_scan(
        T, blockIdx.x, [&](int i) { return o.flag[i]; },
        [&](int i, int v, int excl, int tot) {
            if (v && excl < fp.numFeatures) {
                o.fin_kp[excl] = o.cand[i];
                o.fin_ft[excl] = o.candF[i];
            }
        });
    // the other CTAs of this launch still read counts[2] (T): the new count is parked in counts[3] (free after the
    // refinement stage) and moved by dedup_publish_kernel
    if (blockIdx.x == 0 && threadIdx.x == 0) o.counts[3] = total;
}

__global__ void dedup_publish_kernel(const __grid_constant__ FrameParams fp) {
    const int oi = fp.o_begin + threadIdx.x;
    if (oi < fp.o_end) fp.oct[oi].counts[2] = fp.oct[oi].counts[3];
}

// ---------------------------------------------------------------- descriptor
constexpr int kDescThreads = 128;
constexpr int kDescWarps = kDescThreads / 32;
constexpr int kHistSize = 6 * 6 * 10;
constexpr int kPrivSlots = 32;  // 2 x 2 spatial neighbours x 8 orientation bins (bin 8 is folded onto 0 here, not in the epilogue)

// The 4x4 sample cells (rbin_int, cbin_int) of the descriptor window, packed as R*4+C. Row and column 0 cover rbin in
// (-1,1) (truncating modff, quirk Q13), so cell (0,0) is four units of area, the other row-0 / column-0 cells two, the
// rest one: 25 units, dealt to the four warps as 6 + 6 + 6 + 7. (A CTA-wide queue popped with a shared-memory atomic
// balanced slightly better but cost ~25 instructions per pop: ptxas wraps the atomic into its warp-aggregation sequence.)
__constant__ signed char kWarpCells[4][6] = {{0, 1, -1, -1, -1, -1}, {2, 3, 5, 6, -1, -1}, {4, 8, 7, 9, -1, -1}, {12, 10, 11, 13, 14, 15}};
static_assert(kDescWarps == 4, "kWarpCells deals the cells to four warps");
constexpr int kDescCtasPerSm = 8;  // persistent CTAs per SM (61 registers x 128 threads, 22 KB of shared memory each)
constexpr int kRing = 64;  // per-warp ring of accepted samples (at most 31 pending + 32 appended)

// Everything a warp needs to process one cell of the window, computed once per keypoint by one thread per cell
// (three LDS.128 per cell instead of ~70 warp instructions of setup in each of the warps that own cells).
struct __align__(16) CellInfo {
    int i0, j0, wj, n;            // bounding box origin (patch offsets), its width, its number of points (0: empty)
    int step_r, step_c, hbase;    // the (row, column) advance of 32 consecutive box points; histogram index of the cell's slot 0
    float inv_wj;                 // 1 / wj (for the lanes' first points)
    float r_gt, r_lt, c_gt, c_lt; // truncf(rbin) == R, truncf(cbin) == C written as open intervals
    float fR, fC;
    int pad0, pad1;
};

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {  // x must be a normal number
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rsqrt_approx(float x) {  // x must be a normal number
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// obin of one descriptor sample exactly as the reference computes it (fast_atan2f, SiftOps.cu:403-419, and
// `fast_atan2f(dy, dx) / M_PI * 180.f`, `(grad_ori - angle) * bins_per_degree`, :565-568), in the instruction
// sequence of its sm_100a SASS: r by IEEE division, FMUL(r, 0.1963) -> FFMA(., r, -0.9817) -> FFMA(., r, base),
// the degree conversion in double. Out of line: it runs for a few samples in 10,000 and must not cost the vote
// loop registers.
__device__ __noinline__ float exact_obin(float dx, float dy, float abs_y, bool neg, float angle) {
    const float r = neg ? __fdiv_rn(__fadd_rn(dx, abs_y), __fsub_rn(abs_y, dx)) : __fdiv_rn(__fsub_rn(dx, abs_y), __fadd_rn(dx, abs_y));
    float a = __fmaf_rn(__fmaf_rn(__fmul_rn(r, 0.1963f), r, -0.9817f), r, neg ? 2.356194490f : 0.785398163f);
    a = dy < 0.f ? -a : a;
    float grad_ori = (float)((double)a / 3.14159265358979323846 * 180.0);
    if (grad_ori < 0.f) grad_ori = __fadd_rn(grad_ori, 360.f);
    return __fmul_rn(__fsub_rn(grad_ori, angle), 8 / 360.f);
}

// One CTA per returned keypoint (global index g over octaves in ascending order, truncated at
// numFeatures: utils/CudaMemcpyUtils.cu:38-67). A launch covers an octave range, so that the descriptors of
// octave 0 can start while the small octaves are still being detected; the range that ends at the last
// octave publishes the frame's keypoint count.
//
// The reference (SiftOps.cu:448-604) walks the (2*radius+1)^2 bounding patch and fires 8 shared-memory
// float atomics per sample that lands in the rotated 5x5-cell window. Shared atomics retire about one lane
// per cycle per SM, which bounded the first version of this kernel. Here every warp OWNS whole cells of
// the window (pulled from a CTA-wide queue, largest first):
//   scan   the warp walks the bounding box of its cell, 32 points per step, applies the reference's
//          acceptance test restricted to the cell and appends the accepted (i, j) to a warp-private ring
//          (ballot + popc, no atomics);
//   vote   whenever 32 accepted samples are pending, all 32 lanes process one each (gradient, weight,
//          trilinear split). They all vote into the same 2x2x8 neighbourhood, so each lane accumulates
//          into a private bank-conflict-free column of shared memory with plain LDS/FFMA/STS;
//   flush  at the end of a cell lane s sums slot s over the 32 columns (skewed, conflict-free, fixed
//          order), converts to 2^-15 fixed point and adds it to the CTA histogram with one integer
//          atomic: integer sums are order-independent, so the descriptor is bit-reproducible run to run.
// The set of samples and every per-sample coordinate (c_rot, r_rot, bins, fractions) are computed exactly
// as in the reference; magnitude, angle and weight use the MUFU approximations (a few ulp; the result is
// rounded to fp16).
__global__ void __launch_bounds__(kDescThreads, kDescCtasPerSm)
descriptor_kernel(const __grid_constant__ FrameParams fp) {
    // this launch covers octaves [o_begin, o_end); `first` keypoints of lower octaves precede it in the output
    int first = 0;
    for (int i = 0; i < fp.o_begin; i++) first += fp.oct[i].counts[2];
    int range_total = 0;
    for (int i = fp.o_begin; i < fp.o_end; i++) range_total += fp.oct[i].counts[2];
    if (blockIdx.x == 0 && threadIdx.x == 0 && fp.o_end == fp.nOct) *fp.total = min(first + range_total, fp.numFeatures);
    const int limit = min(range_total, fp.numFeatures - first);  // keypoints of this range that are returned
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    __shared__ int s_hist[kHistSize];
    __shared__ float s_f[kHistSize];
    __shared__ float s_desc[128];
    __shared__ float s_norm;
    __shared__ __align__(128) float s_priv[kDescWarps][kPrivSlots][32];
    __shared__ int s_ring[kDescWarps][kRing];
    __shared__ __align__(16) CellInfo s_cell[16];
    float* priv = &s_priv[warp][0][lane];
#pragma unroll
    for (int k = 0; k < kPrivSlots; k++) priv[k * 32] = 0.f;  // every flush leaves the columns zeroed again

    for (int b = blockIdx.x; b < limit; b += gridDim.x) {  // persistent CTAs, grid-stride over the keypoints
    int oi = fp.o_begin, j = b;
    while (j >= fp.oct[oi].counts[2]) { j -= fp.oct[oi].counts[2]; oi++; }
    const int g = first + b;
    const OctParams& o = fp.oct[oi];
    const float3 kp = o.fin_kp[j];
    const float4 ft = o.fin_ft[j];
    for (int i = tid; i < kHistSize; i += kDescThreads) s_hist[i] = 0;  // the barrier after the cell boxes below covers these

    // unpackOctave (SiftOps.cu:389-401)
    const int packed = (int)ft.x;
    int octv = packed & 255;
    const int layer = (packed >> 8) & 255;
    octv = octv < 128 ? octv : (-128 | octv);
    const float scale = octv >= 0 ? __fdiv_rn(1.f, (float)(1 << octv)) : (float)(1 << -octv);
    const int hdim = o.pitch;
    const int ow = o.w, oh = o.h;
    const long stack_len = (long)o.pitch * o.h * fp.nLayers;
    const long img_off = (long)((float)hdim * (float)o.h * (float)layer);
    const int ix = __float2int_rn(__fmul_rn(kp.x, scale));
    const int iy = __float2int_rn(__fmul_rn(kp.y, scale));
    float angle = __fsub_rn(360.f, ft.w);
    if (fabsf(__fsub_rn(angle, 360.f)) < 1e-6f) angle = 0.f;
    const float scl = __fmul_rn(__fmul_rn(ft.y, scale), 0.5f);
    constexpr float deg2rad = (float)(3.14159265358979323846 / 180.0);
    float cos_u, sin_u;
    __sincosf(__fmul_rn(angle, deg2rad), &sin_u, &cos_u);
    constexpr float bins_per_degree = 8 / 360.f;
    const float hist_width = __fmul_rn(3.f, scl);
    int radius = __float2int_rn(__fmul_rn(__fmul_rn(__fmul_rn(hist_width, 1.4142135623730951f), 5.f), 0.5f));
    const float diag = sqrtf(__fmaf_rn((float)o.w, (float)o.w, __fmul_rn((float)o.h, (float)o.h)));
    radius = (int)fmin((double)radius, (double)diag);
    const float cos_t = __fdiv_rn(cos_u, hist_width);
    const float sin_t = __fdiv_rn(sin_u, hist_width);
    const float* stack = o.gauss + img_off;
    const int stack_left = (int)(stack_len - img_off);  // floats from this layer to the end of the octave's stack
    const float ch = cos_u * hist_width, sh = sin_u * hist_width;  // window -> patch coordinates

    if (tid < 16) {
        // cell tid = R*4 + C: its extent in rotated coordinates (r_rot = rbin - 1.5), then its bounding box in patch
        // offsets: i = (r cos - c sin) * hist_width, j = (r sin + c cos) * hist_width. floor/ceil leave up to a pixel of
        // slack, six orders of magnitude more than the rounding of these products. One thread per cell, once per
        // keypoint (the warps that own the cells used to recompute this, 83 warp instructions per cell).
        const int R = tid >> 2, C = tid & 3;
        const float fR = (float)R, fC = (float)C;
        const float rlo = (R == 0 ? -1.f : fR) - 1.5f, rhi = fR - 0.5f;
        const float clo = (C == 0 ? -1.f : fC) - 1.5f, chi = fC - 0.5f;
        const float ia = rlo * ch, ib = rhi * ch, ic = -clo * sh, id = -chi * sh;
        const float ja = rlo * sh, jb = rhi * sh, jc = clo * ch, jd = chi * ch;
        // ... clipped to the patch (|i|, |j| <= radius) and to the pixels the reference accepts (0 < x < w, 0 < y < h
        // strictly, SiftOps.cu:541-547): the scan then needs no per-point image test
        const int i0 = max(max(-radius, 1 - iy), __float2int_rd(fminf(ia, ib) + fminf(ic, id)));
        const int i1 = min(min(radius, oh - 1 - iy), __float2int_ru(fmaxf(ia, ib) + fmaxf(ic, id)));
        const int j0 = max(max(-radius, 1 - ix), __float2int_rd(fminf(ja, jb) + fminf(jc, jd)));
        const int j1 = min(min(radius, ow - 1 - ix), __float2int_ru(fmaxf(ja, jb) + fmaxf(jc, jd)));
        const int wj = j1 - j0 + 1, wi = i1 - i0 + 1;
        const bool empty = wj <= 0 || wi <= 0;
        const float inv_wj = __frcp_rn((float)max(wj, 1));
        // (k + 0.5) / wj is never within rounding distance of an integer, so this is an exact floor
        const int step_r = empty ? 0 : __float2int_rd(32.5f * inv_wj);
        CellInfo ci;
        ci.i0 = i0; ci.j0 = j0; ci.wj = wj; ci.n = empty ? 0 : wi * wj;
        ci.step_r = step_r; ci.step_c = 32 - step_r * wj;
        ci.hbase = ((R + 1) * 6 + (C + 1)) * 10;
        ci.inv_wj = inv_wj;
        // truncf(rbin) == R written as an interval: (-1, 1) for R == 0, [R, R+1) otherwise
        ci.r_gt = R == 0 ? -1.f : __uint_as_float(__float_as_uint(fR) - 1u); ci.r_lt = fR + 1.f;
        ci.c_gt = C == 0 ? -1.f : __uint_as_float(__float_as_uint(fC) - 1u); ci.c_lt = fC + 1.f;
        ci.fR = fR; ci.fC = fC;
        ci.pad0 = ci.pad1 = 0;
        s_cell[tid] = ci;
    }
    __syncthreads();

    int* ring = s_ring[warp];
    const unsigned lt_mask = (1u << lane) - 1u;
    const int lane_hoff = ((lane >> 4) * 6 + ((lane >> 3) & 1)) * 10 + (lane & 7);  // slot `lane` = (dr*2 + dc)*8 + bin of a cell
    // gradient reads: one 64-bit base per keypoint (the patch centre), 32-bit element offsets per sample
    const int a_centre = iy * hdim + ix;
    const float* centre = stack + a_centre;
    const int off_lim = stack_left - hdim - a_centre;  // the row below exists in the stack iff off < off_lim
    for (int qi = 0; qi < 6; qi++) {
        // the broadcast tells the compiler that t (and with it the whole cell state) is warp-uniform: uniform datapath
        const int t = __shfl_sync(0xffffffffu, (int)kWarpCells[warp][qi], 0);
        if (t < 0) break;
        const int4 c0 = reinterpret_cast<const int4*>(&s_cell[t])[0];
        const int i0 = c0.x, j0 = c0.y, wj = c0.z, n = c0.w;
        if (n <= 0) continue;
        const int4 c1 = reinterpret_cast<const int4*>(&s_cell[t])[1];
        const float4 c2 = reinterpret_cast<const float4*>(&s_cell[t])[2];
        const float2 c3 = reinterpret_cast<const float2*>(&s_cell[t])[6];
        const int step_r = c1.x, step_c = c1.y, hbase = c1.z;
        const float r_gt = c2.x, r_lt = c2.y, c_gt = c2.z, c_lt = c2.w;
        const float fR = c3.x, fC = c3.y;
        // (k + 0.5) / wj is never within rounding distance of an integer, so this is an exact floor
        const int prow0 = __float2int_rd(((float)lane + 0.5f) * __int_as_float(c1.w));
        int pi = i0 + prow0, pj = j0 + (lane - prow0 * wj);  // this lane's point of the current 32-point step
        const int jend = j0 + wj;
        int rem = n - lane;  // > 0 while this lane's point is inside the box
        int head = 0, tail = 0;  // accepted / consumed sample counts of this cell (warp-uniform)
        int scanned = 0;
        for (;;) {
            // ---- scan: until a full batch is pending or the box is exhausted
            while (scanned < n && head - tail < 32) {
                const int idx_i = pi, idx_j = pj;
                const bool in_box = rem > 0;
                scanned += 32;
                rem -= 32;
                pi += step_r;
                pj += step_c;
                if (pj >= jend) { pj -= wj; pi++; }
                const float fi = (float)idx_i, fj = (float)idx_j;
                const float c_rot = __fmaf_rn(fj, cos_t, -__fmul_rn(fi, sin_t));
                const float r_rot = __fmaf_rn(fi, cos_t, __fmul_rn(fj, sin_t));
                const float cbin = __fadd_rn(c_rot, 1.5f), rbin = __fadd_rn(r_rot, 1.5f);
                // the reference's acceptance (rbin, cbin in (-1,4); the pixel test is in the box), restricted to this cell
                const bool ok = in_box && rbin > r_gt && rbin < r_lt && cbin > c_gt && cbin < c_lt;
                const unsigned b = __ballot_sync(0xffffffffu, ok);
                if (ok) ring[(head + __popc(b & lt_mask)) & (kRing - 1)] = (idx_i & 0xffff) | (idx_j << 16);
                head += __popc(b);
            }
            const int pend = head - tail;
            if (pend == 0) break;
            __syncwarp();
            // ---- vote: one accepted sample per lane
            if (lane < pend) {
                const int e = ring[(tail + lane) & (kRing - 1)];
                const int idx_i = (int)(short)(e & 0xffff), idx_j = e >> 16;
                const float fi = (float)idx_i, fj = (float)idx_j;
                const float c_rot = __fmaf_rn(fj, cos_t, -__fmul_rn(fi, sin_t));
                const float r_rot = __fmaf_rn(fi, cos_t, __fmul_rn(fj, sin_t));
                const float cbin = __fadd_rn(c_rot, 1.5f), rbin = __fadd_rn(r_rot, 1.5f);
                // quirk Q12: px == w-1 / py == h-1 are accepted, so these reads may land in the row
                // padding, the next row or the next layer; clamp only at the end of the stack.
                const int off = idx_i * hdim + idx_j;
                const float *pc, *pu, *pd;  // one IMAD.WIDE each
                asm("mad.wide.s32 %0, %1, 4, %2;" : "=l"(pc) : "r"(off), "l"(centre));
                asm("mad.wide.s32 %0, %1, 4, %2;" : "=l"(pu) : "r"(off - hdim), "l"(centre));
                asm("mad.wide.s32 %0, %1, 4, %2;" : "=l"(pd) : "r"(off + hdim), "l"(centre));
                const float v_l = __ldg(pc - 1), v_r = __ldg(pc + 1);
                const float v_up = __ldg(pu);
                const float v_dn = off < off_lim ? __ldg(pd) : 0.f;
                const float dx = __fsub_rn(v_r, v_l);
                const float dy = __fsub_rn(v_up, v_dn);
                // rbin_int = truncf(rbin) is R (as -0.0 for rbin in (-1,0): the fraction is then negative)
                const float rbin_frac = __fsub_rn(rbin, fR), cbin_frac = __fsub_rn(cbin, fC);
                const float d2 = __fmaf_rn(c_rot, c_rot, __fmul_rn(r_rot, r_rot));
                // Gaussian weight x 2^15 (the fixed-point scale of the integer sum, undone after it) in one MUFU.EX2
                const float weight = ex2_approx(__fmaf_rn(d2, -0.125f * 1.4426950408889634f, 15.f));
                const float g2 = __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
                const float grad_mag = g2 * rsqrt_approx(fmaxf(g2, 1e-30f));
                // fast_atan2f (SiftOps.cu:403-419) with one approximate division
                const float abs_y = fabsf(dy) + 1e-10f;
                const bool neg = dx < 0.f;
                const float num = neg ? dx + abs_y : dx - abs_y;
                const float den = neg ? abs_y - dx : dx + abs_y;
                const float q = num * rcp_approx(den);  // den >= 1e-10
                float ang = __fmaf_rn(__fmaf_rn(0.1963f * q, q, -0.9817f), q, neg ? 2.356194490f : 0.785398163f);
                ang = dy < 0.f ? -ang : ang;
                float grad_ori = ang * 57.29577951308232f;
                grad_ori = grad_ori < 0.f ? grad_ori + 360.f : grad_ori;
                float obin = __fmul_rn(__fsub_rn(grad_ori, angle), bins_per_degree);
                float obin_t = truncf(obin);
                float obin_frac = __fsub_rn(obin, obin_t);
                // The vote is a discontinuous function of the angle in two places: where fast_atan2f changes sign (the
                // +360 wrap meets the truncating modff: obin and obin - 8 do not vote alike, quirk Q11) and where obin
                // crosses a (negative) integer. The approximations above are good to ~2e-6 bins, so within 1e-4 bins /
                // 1e-5 rad of either the reference's own instruction sequence decides (SASS of SiftOps.cu:403-419,565:
                // IEEE division, FMUL + 2 FFMA, then double / pi * 180): identical bins, ~2 samples in 10,000.
                // |frac| < 1e-4 or > 1 - 1e-4  <=>  ||frac| - 0.5| > 0.4999
                if (fabsf(ang) < 1e-5f || fabsf(fabsf(obin_frac) - 0.5f) > 0.4999f) {
                    obin = exact_obin(dx, dy, abs_y, neg, angle);
                    obin_t = truncf(obin);
                    obin_frac = __fsub_rn(obin, obin_t);
                }
                const float mag = grad_mag * weight;
                const int ob = (int)obin_t & 7;  // -7..8 -> 0..7 (obin_int < 0: += 8, >= 8: -= 8)
                const float v_r1 = mag * rbin_frac, v_r0 = mag * (1.f - rbin_frac);
                const float omc = 1.f - cbin_frac, omo = 1.f - obin_frac;
                const float v11 = v_r1 * cbin_frac, v10 = v_r1 * omc;
                const float v01 = v_r0 * cbin_frac, v00 = v_r0 * omc;
                float* p = priv + ob * 32;                  // slot (dr*2 + dc)*8 + ob of this lane's column
                float* p1 = priv + ((ob + 1) & 7) * 32;     // the next orientation bin, 8 wraps to 0
                float a0 = p[0], a1 = p1[0], a2 = p[8 * 32], a3 = p1[8 * 32];
                float a4 = p[16 * 32], a5 = p1[16 * 32], a6 = p[24 * 32], a7 = p1[24 * 32];
                a0 = __fmaf_rn(v00, omo, a0); a1 = __fmaf_rn(v00, obin_frac, a1);
                a2 = __fmaf_rn(v01, omo, a2); a3 = __fmaf_rn(v01, obin_frac, a3);
                a4 = __fmaf_rn(v10, omo, a4); a5 = __fmaf_rn(v10, obin_frac, a5);
                a6 = __fmaf_rn(v11, omo, a6); a7 = __fmaf_rn(v11, obin_frac, a7);
                p[0] = a0; p1[0] = a1; p[8 * 32] = a2; p1[8 * 32] = a3;
                p[16 * 32] = a4; p1[16 * 32] = a5; p[24 * 32] = a6; p1[24 * 32] = a7;
            }
            tail += min(pend, 32);
            __syncwarp();
        }
        // ---- flush: lane s owns slot s = (dr*2 + dc)*8 + bin, i.e. row s of the warp's [slot][lane] block: it sums
        // the row's 32 lane-private entries with eight LDS.128 (skewed by s: conflict-free within a quarter warp,
        // fixed order) and leaves the row zeroed for the next cell
        float4* row4 = reinterpret_cast<float4*>(&s_priv[warp][lane][0]);
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            float4* p4 = row4 + (k ^ (lane & 7));
            const float4 v = *p4;
            *p4 = make_float4(0.f, 0.f, 0.f, 0.f);
            acc += v.x; acc += v.y; acc += v.z; acc += v.w;
        }
        const int tot = __float2int_rn(acc);
        if (tot != 0) atomicAdd(s_hist + hbase + lane_hoff, tot);
        __syncwarp();
    }
    __syncthreads();
    for (int i = tid; i < kHistSize; i += kDescThreads) s_f[i] = __fmul_rn((float)s_hist[i], kUnfix32);
    __syncthreads();
    // copyToDescriptor (SiftOps.cu:431-446): fold orientation bins 8,9 onto 0,1
    {
        const int cell = tid >> 3, k = tid & 7;
        const int ci = cell >> 2, cj = cell & 3;
        const int base = ((ci + 1) * 6 + (cj + 1)) * 10;
        s_desc[tid] = s_f[base + k];  // the wrap of bin 8 onto 0 (SiftOps.cu:431-446) already happened per vote
    }
    __syncthreads();
    if (tid == 0) {
        // norm / clip_lower / norm, sequential FMA chains as the reference's lane 0 (SiftOps.cu:606-615)
        float cnt = 0.f;
        for (int i = 0; i < 128; i++) cnt = __fmaf_rn(s_desc[i], s_desc[i], cnt);
        const float nrm = sqrtf(cnt);
        const float thr = (float)((double)nrm * 0.2);
        cnt = 0.f;
        for (int i = 0; i < 128; i++) {
            const float v = fminf(s_desc[i], thr);
            s_desc[i] = v;
            cnt = __fmaf_rn(v, v, cnt);
        }
        s_norm = fmaxf(sqrtf(cnt), 1e-7f);
    }
    __syncthreads();
    const float v = __saturatef(__fdiv_rn(s_desc[tid], s_norm));
    fp.out_desc[(size_t)g * 128 + tid] = __float2half(__fmul_rn(v, 512.f));
    if (tid == 0) {
        fp.out_kpts[g] = kp;
        fp.out_feats[g] = ft;
    }
    }  // keypoint loop (the barriers above order the reuse of the shared buffers)
}

}  // namespace

void launch_orientation(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s) {
    long cap = 0;  // CTA per keypoint, grid-stride; small octave ranges cannot fill a 16-deep grid
    for (int i = o_begin; i < o_end; i++) cap += fp.oct[i].cap;
    const int grid = (int)(cap < 148L * 16 ? (cap > 0 ? cap : 1) : 148L * 16);
    orientation_kernel<<<grid, kOriThreads, 0, s>>>(octave_range(fp, o_begin, o_end));
}
void launch_compact_oriented(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s) {
    int maxcap = 1;
    for (int i = o_begin; i < o_end; i++) maxcap = fp.oct[i].cap > maxcap ? fp.oct[i].cap : maxcap;
    dim3 grid((maxcap + kScanBlock - 1) / kScanBlock, o_end - o_begin);
    compact_oriented_kernel<<<grid, kScanBlock, 0, s>>>(octave_range(fp, o_begin, o_end));
}
void launch_dedup(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s) {
    int maxcap = 1;
    for (int i = o_begin; i < o_end; i++) maxcap = fp.oct[i].cap > maxcap ? fp.oct[i].cap : maxcap;
    dim3 grid((maxcap + kScanBlock - 1) / kScanBlock, o_end - o_begin);
    dedup_flag_kernel<<<grid, kScanBlock, 0, s>>>(octave_range(fp, o_begin, o_end));
    dedup_compact_kernel<<<grid, kScanBlock, 0, s>>>(octave_range(fp, o_begin, o_end));
    dedup_publish_kernel<<<1, 32, 0, s>>>(octave_range(fp, o_begin, o_end));
}
static int g_desc_priority = 0;

cudaError_t descriptor_setup() {  // once per process, outside stream capture
    int greatest = 0;
    return cudaDeviceGetStreamPriorityRange(&g_desc_priority, &greatest);
}

cudaError_t launch_descriptor(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s) {
    // upper bound on the keypoints of the range: its slot capacity, and never more than the output holds
    long cap = 0;
    for (int i = o_begin; i < o_end; i++) cap += fp.oct[i].cap;
    long grid = cap < (long)fp.numFeatures ? cap : (long)fp.numFeatures;
    static int per_sm = 0;
    if (!per_sm) {  // development switch for same-box A/B runs: B200SIFT_DESC_CTAS=<persistent CTAs per SM>
        const char* e = std::getenv("B200SIFT_DESC_CTAS");
        per_sm = e ? std::atoi(e) : kDescCtasPerSm;
        if (per_sm < 1 || per_sm > kDescCtasPerSm) per_sm = kDescCtasPerSm;
    }
    static int waves = 0;
    if (!waves) {  // development switch: B200SIFT_DESC_WAVES=k -> up to k waves of shorter-lived CTAs instead of one persistent wave
        const char* e = std::getenv("B200SIFT_DESC_WAVES");
        waves = e ? std::atoi(e) : 1;
        if (waves < 1 || waves > 64) waves = 1;
    }
    if (grid > 148L * per_sm * waves) grid = 148L * per_sm * waves;
    // Lowest stream priority: the descriptor grids would otherwise fill every CTA slot and starve the short
    // latency-bound kernels of the other graph branches (the priority is recorded in the captured node).
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(grid > 0 ? grid : 1));
    cfg.blockDim = dim3(kDescThreads);
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributePriority;
    attr[0].val.priority = g_desc_priority;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, descriptor_kernel, octave_range(fp, o_begin, o_end));
}

}  // namespace b200sift
