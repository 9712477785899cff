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
// (|gradient| < 170, weight <= 1, a few hundred samples per bin). Quantisation: <= 1.5e-5 per vote, i.e.
// ~1e-6 relative on a bin -- below the fp32 rounding noise of the reference's own float accumulation.
constexpr float kFix32 = 32768.f;
constexpr float kUnfix32 = 1.f / 32768.f;
__device__ __forceinline__ void fix_add32(int* slot, float v_scaled) { atomicAdd(slot, __float2int_rn(v_scaled)); }

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
    for (int i = 0; i < fp.nOct; i++) total += fp.oct[i].counts[1];

    for (int g = blockIdx.x; g < total; g += gridDim.x) {
        int oi = 0, k = g;
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
        const float scale = __fdiv_rn(__fmul_rn(1.5f, ft.y), (float)(1 << (oi + 1)));
        const float radius = roundf(__fmul_rn(3.f, scale));
        const int base_px = (int)roundf(__fdiv_rn(kp.x, (float)(1 << oi)));
        const int base_py = (int)roundf(__fdiv_rn(kp.y, (float)(1 << oi)));
        int side = (int)__fmaf_rn(2.f, radius, 1.f);
        if (side > 1024) side = 1024;  // only a corrupt stale slot could ask for more
        const int total_dim = side * side;
        const float weight_factor = __fdiv_rn(-0.5f, __fmul_rn(scale, scale));
        int layer = (int)kp.z;
        layer = min(max(layer, 0), 4);
        const float* g0 = o.gauss + ls * layer;  // DoG[layer] = G[layer+1] - G[layer]
        const float fh1 = (float)o.h - 1.f, fw1 = (float)o.w - 1.f;

        auto geometry = [&](int idx, int& px_off, int& py_off, const float*& p) -> bool {
            if (idx >= total_dim) return false;
            py_off = (int)__fsub_rn((float)(idx / side), radius);
            const int py = base_py + py_off;
            px_off = (int)__fsub_rn((float)(idx % side), radius);
            const int px = base_px + px_off;
            if (py < 1 || (float)py >= fh1 || px < 1 || (float)px >= fw1) return false;
            p = g0 + (size_t)py * h_dim + px;
            return true;
        };
        auto vote = [&](const float v[8], int px_off, int py_off) -> OriSample {
            const float dx = __fsub_rn(__fsub_rn(v[0], v[1]), __fsub_rn(v[2], v[3]));
            const float dy = __fsub_rn(__fsub_rn(v[4], v[5]), __fsub_rn(v[6], v[7]));
            const float mag = __fsqrt_rd(__fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
            float ori = __fmul_rn(__fdiv_rn(atan2f(dy, dx), 3.14159265358979323846f), 180.f);
            if (ori < 0.f) ori = __fadd_rn(ori, 360.f);
            const float weight = __expf(__fmul_rn(weight_factor, (float)(px_off * px_off + py_off * py_off)));
            int bin = __float2int_rd(__fmul_rn(__fdiv_rn(ori, 360.f), 36.f)) % kOriBins;
            if (bin < 0) bin += kOriBins;  // NaN guard; never taken for finite input
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

// One CTA per octave: prefix over per-keypoint peak counts -> cand/candF in (k, bin) order; the
// entry whose prefix equals the total T is not written (quirk Q5 again).
__global__ void __launch_bounds__(1024)
compact_oriented_kernel(const __grid_constant__ FrameParams fp) {
    const OctParams& o = fp.oct[blockIdx.x];
    const int M = o.counts[1];
    __shared__ int warp_sums[32];
    __shared__ int s_total, carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int local = 0;
    for (int i = threadIdx.x; i < M; i += 1024) local += o.oricnt[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) local += __shfl_xor_sync(0xffffffffu, local, d);
    if (lane == 0) warp_sums[warp] = local;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tsum = 0;
        for (int i = 0; i < 32; i++) tsum += warp_sums[i];
        s_total = tsum;
        carry = 0;
        o.counts[2] = tsum;  // T
    }
    __syncthreads();
    const int total = s_total;
    for (int base = 0; base < M; base += 1024) {
        const int k = base + threadIdx.x;
        const int v = k < M ? o.oricnt[k] : 0;
        int incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int u = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += u;
        }
        __syncthreads();
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int ws = warp_sums[lane];
            int wi = ws;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int u = __shfl_up_sync(0xffffffffu, wi, d);
                if (lane >= d) wi += u;
            }
            warp_sums[lane] = wi - ws;
        }
        __syncthreads();
        const int c0 = carry;
        const int excl = c0 + warp_sums[warp] + incl - v;
        for (int c = 0; c < v; c++) {
            const int p = excl + c + 1;
            if (p < total) {
                const int slot = M + k * kOriBins + c;
                o.cand[p - 1] = o.coll[slot];
                o.candF[p - 1] = o.collF[slot];
            }
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
}

// ---------------------------------------------------------------- descriptor
// fast_atan2f (SiftOps.cu:403-419), contraction as in the reference SASS.
__device__ __forceinline__ float fast_atan2f(float y, float x) {
    const float abs_y = __fadd_rn(fabsf(y), 1e-10f);
    float r, angle;
    if (x < 0.0f) {
        r = __fdiv_rn(__fadd_rn(x, abs_y), __fsub_rn(abs_y, x));
        angle = 2.356194490f;
    } else {
        r = __fdiv_rn(__fsub_rn(x, abs_y), __fadd_rn(x, abs_y));
        angle = 0.785398163f;
    }
    angle = __fmaf_rn(__fmaf_rn(__fmul_rn(0.1963f, r), r, -0.9817f), r, angle);
    return (y < 0.0f) ? -angle : angle;
}

constexpr int kDescThreads = 128;
constexpr int kHistSize = 6 * 6 * 10;

// One CTA per returned keypoint (global index g over octaves in ascending order, truncated at
// numFeatures: utils/CudaMemcpyUtils.cu:38-67). 128 threads stride over the rotated patch.
__global__ void __launch_bounds__(kDescThreads)
descriptor_kernel(const __grid_constant__ FrameParams fp) {
    const int g = blockIdx.x;
    int oi = 0, j = g, sum = 0;
    bool found = false;
    for (int i = 0; i < fp.nOct; i++) {
        const int T = fp.oct[i].counts[2];
        if (!found && j < T) { oi = i; found = true; }
        if (!found) j -= T;
        sum += T;
    }
    if (g == 0 && threadIdx.x == 0) *fp.total = min(sum, fp.numFeatures);
    if (!found) return;
    const OctParams& o = fp.oct[oi];
    const float3 kp = o.cand[j];
    const float4 ft = o.candF[j];
    const int tid = threadIdx.x;

    __shared__ int s_hist[kHistSize];
    __shared__ float s_f[kHistSize];
    __shared__ float s_desc[128];
    __shared__ float s_norm;
    for (int i = tid; i < kHistSize; i += kDescThreads) s_hist[i] = 0;
    __syncthreads();

    // unpackOctave (SiftOps.cu:389-401)
    const int packed = (int)ft.x;
    int octv = packed & 255;
    const int layer = (packed >> 8) & 255;
    octv = octv < 128 ? octv : (-128 | octv);
    const float scale = octv >= 0 ? __fdiv_rn(1.f, (float)(1 << octv)) : (float)(1 << -octv);
    const int hdim = o.pitch;
    const long stack_len = (long)o.pitch * o.h * kGaussLayers;
    const long img_off = (long)((float)hdim * (float)o.h * (float)layer);
    const float x = (float)__float2int_rn(__fmul_rn(kp.x, scale));
    const float y = (float)__float2int_rn(__fmul_rn(kp.y, scale));
    float angle = __fsub_rn(360.f, ft.w);
    if (fabsf(__fsub_rn(angle, 360.f)) < 1e-6f) angle = 0.f;
    const float scl = __fmul_rn(__fmul_rn(ft.y, scale), 0.5f);
    constexpr float deg2rad = (float)(3.14159265358979323846 / 180.0);
    float cos_t, sin_t;
    __sincosf(__fmul_rn(angle, deg2rad), &sin_t, &cos_t);
    constexpr float bins_per_degree = 8 / 360.f;
    const float hist_width = __fmul_rn(3.f, scl);
    int radius = __float2int_rn(__fmul_rn(__fmul_rn(__fmul_rn(hist_width, 1.4142135623730951f), 5.f), 0.5f));
    const float diag = sqrtf(__fmaf_rn((float)o.w, (float)o.w, __fmul_rn((float)o.h, (float)o.h)));
    radius = (int)fmin((double)radius, (double)diag);
    cos_t = __fdiv_rn(cos_t, hist_width);
    sin_t = __fdiv_rn(sin_t, hist_width);
    const int side = 2 * radius + 1;
    const int total_dim = side * side;
    const float fw = (float)o.w, fh = (float)o.h;
    const float* stack = o.gauss;

    // idx = tid + k*128 walks the patch row-major; (row, col) advance incrementally instead of idx / side, idx % side
    const int step_r = kDescThreads / side, step_c = kDescThreads % side;
    int prow = tid / side, pcol = tid % side;
    for (int idx = tid; idx < total_dim; idx += kDescThreads) {
        const int idx_i = prow - radius, idx_j = pcol - radius;
        prow += step_r;
        pcol += step_c;
        if (pcol >= side) { pcol -= side; prow++; }
        const float fi = (float)idx_i, fj = (float)idx_j;
        const float c_rot = __fmaf_rn(fj, cos_t, -__fmul_rn(fi, sin_t));
        const float r_rot = __fmaf_rn(fi, cos_t, __fmul_rn(fj, sin_t));
        const float cbin = __fadd_rn(c_rot, 1.5f), rbin = __fadd_rn(r_rot, 1.5f);
        const int base_py = (int)__fadd_rn(y, fi), base_px = (int)__fadd_rn(x, fj);
        if (rbin > -1.f && rbin < 4.f && cbin > -1.f && cbin < 4.f && (float)base_px > 0.f && (float)base_px < fw &&
            (float)base_py > 0.f && (float)base_py < fh) {
            // quirk Q12: px == w-1 / py == h-1 pass the test above, so these reads may land in the
            // row padding, the next row or the next layer; clamp only at the end of the stack.
            const long a = img_off + (long)base_py * hdim + base_px;
            const long a_dn = a + hdim;
            const float v_l = __ldg(stack + a - 1), v_r = __ldg(stack + a + 1);
            const float v_up = __ldg(stack + a - hdim);
            const float v_dn = a_dn < stack_len ? __ldg(stack + a_dn) : 0.f;
            const float dx = __fsub_rn(v_r, v_l);
            const float dy = __fsub_rn(v_up, v_dn);
            const float rbin_int = truncf(rbin), cbin_int = truncf(cbin);
            const float rbin_frac = __fsub_rn(rbin, rbin_int), cbin_frac = __fsub_rn(cbin, cbin_int);
            const float weight = __expf(__fmul_rn(__fmaf_rn(c_rot, c_rot, __fmul_rn(r_rot, r_rot)), -0.125f));
            const float grad_mag = __fsqrt_rn(__fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
            float grad_ori = (float)(((double)fast_atan2f(dy, dx) / 3.14159265358979323846) * 180.0);
            grad_ori = grad_ori < 0.f ? __fadd_rn(grad_ori, 360.f) : grad_ori;
            const float obin = __fmul_rn(__fsub_rn(grad_ori, angle), bins_per_degree);
            const float mag = __fmul_rn(__fmul_rn(grad_mag, weight), kFix32);  // x 2^15: exact, commutes with the roundings below
            float obin_int = truncf(obin);
            const float obin_frac = __fsub_rn(obin, obin_int);
            if (obin_int < 0.f) obin_int += 8.f;
            if (obin_int >= 8.f) obin_int -= 8.f;
            const int hi = (int)(((rbin_int + 1.f) * 6.f + cbin_int + 1.f) * 10.f + obin_int);
            const float v_r1 = __fmul_rn(mag, rbin_frac), v_r0 = __fmul_rn(mag, __fsub_rn(1.f, rbin_frac));
            const float omc = __fsub_rn(1.f, cbin_frac), omo = __fsub_rn(1.f, obin_frac);
            const float v_rc11 = __fmul_rn(v_r1, cbin_frac), v_rc10 = __fmul_rn(v_r1, omc);
            const float v_rc01 = __fmul_rn(v_r0, cbin_frac), v_rc00 = __fmul_rn(v_r0, omc);
            if (hi >= 0 && hi + 71 < kHistSize) {
                fix_add32(s_hist + hi + 70, __fmul_rn(v_rc11, omo));
                fix_add32(s_hist + hi + 71, __fmul_rn(v_rc11, obin_frac));
                fix_add32(s_hist + hi + 60, __fmul_rn(v_rc10, omo));
                fix_add32(s_hist + hi + 61, __fmul_rn(v_rc10, obin_frac));
                fix_add32(s_hist + hi + 10, __fmul_rn(v_rc01, omo));
                fix_add32(s_hist + hi + 11, __fmul_rn(v_rc01, obin_frac));
                fix_add32(s_hist + hi, __fmul_rn(v_rc00, omo));
                fix_add32(s_hist + hi + 1, __fmul_rn(v_rc00, obin_frac));
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < kHistSize; i += kDescThreads) s_f[i] = __fmul_rn((float)s_hist[i], kUnfix32);
    __syncthreads();
    // copyToDescriptor (SiftOps.cu:431-446): fold orientation bins 8,9 onto 0,1
    {
        const int cell = tid >> 3, k = tid & 7;
        const int ci = cell >> 2, cj = cell & 3;
        const int base = ((ci + 1) * 6 + (cj + 1)) * 10;
        float v = s_f[base + k];
        if (k < 2) v = __fadd_rn(v, s_f[base + 8 + k]);
        s_desc[tid] = v;
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
}

}  // namespace

void launch_orientation(const FrameParams& fp, cudaStream_t s) {
    orientation_kernel<<<148 * 16, kOriThreads, 0, s>>>(fp);
}
void launch_compact_oriented(const FrameParams& fp, cudaStream_t s) { compact_oriented_kernel<<<fp.nOct, 1024, 0, s>>>(fp); }
void launch_descriptor(const FrameParams& fp, cudaStream_t s) {
    descriptor_kernel<<<fp.numFeatures, kDescThreads, 0, s>>>(fp);
}

}  // namespace b200sift
