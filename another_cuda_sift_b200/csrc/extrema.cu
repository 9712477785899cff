// DoG + 3x3x3 extrema + ordered compaction + Taylor refinement.
//
// Replaces image_func/MatOps.cu (minus, findPeaksBlockFirstPass, cub scan over an int-per-pixel
// mask, findPeaksBlockSecondPass), sift_func/SiftOps.cu:63-235 (adjustExtrema, collectKpts) and
// the memsetMask launches of interface/Detector.cu:330-343.
//
//  detect  : one pass over the 6 Gaussian layers of every octave. DoG is formed in registers
//            (G[l+1]-G[l] is a single exact FADD, so nothing is lost by not storing it); a warp
//            slides down a 32-pixel-wide column strip keeping a 3-row window of horizontal
//            3-max / 3-min per DoG layer, so each Gaussian pixel is loaded once per strip.
//            Output: 1 bit per pixel per scale (warp ballot) + per-row popcounts.
//  scan    : exclusive prefix over (scale,row) counts -> raster order (scale, y, x) exactly as
//            the reference's global inclusive scan gives it.
//  scatter : warp per non-empty row, popc/prefix-sum append of (x, y, scale+1) at slot prefix-1
//            while prefix < cap (MatOps.cu:172-179).
//  refine  : thread per candidate, same expression tree and FMA contraction as the reference
//            kernel compiled for sm_100a (checked against its SASS), DoG recomputed on the fly.
//
// Roofline: HBM-bound. Algorithmic bytes of detect = 24 B per pixel of every octave (6 layers
// read once) + 3/8 B per pixel of mask.
#include "common.cuh"

namespace b200sift {

namespace {

constexpr int kDetThreads = 128;

// Raw Gaussian values of one row of a 64-pixel strip: lane l owns pixels 2l, 2l+1 of the strip (one 8-byte
// load per layer); lanes 0 and 31 also fetch the pixel left / right of the strip.
template <int NG>
struct RawRow {
    float2 g[NG];
    float gh[NG];
};

template <int NG>
__device__ __forceinline__ void load_raw(const float* __restrict__ base, size_t ls, int x, int xh, bool halo_lane, RawRow<NG>& r) {
#pragma unroll
    for (int l = 0; l < NG; l++) r.g[l] = __ldg(reinterpret_cast<const float2*>(base + ls * l + x));
#pragma unroll
    for (int l = 0; l < NG; l++) r.gh[l] = halo_lane ? __ldg(base + ls * l + xh) : 0.f;
}

// Warp task = 64-pixel-wide column strip x kDetRows rows of one octave, for NS consecutive scales starting at scale
// s0 (NS + 2 DoG layers from NS + 3 Gaussian layers). Per DoG layer the warp keeps, for the two pixels of every lane,
// the horizontal 3-max / 3-min of the previous row and of the previous two rows combined, so the 3x3 in-plane extrema
// of the centre row cost two max and two min per new row; the raw loads of the next row are issued before the current
// row is processed.
// Mask layout: word 2c of a row holds the even pixels of strip c (bit l <-> x = 64c + 2l), word 2c+1 the
// odd ones; scatter_kernel merges the two in x order.
template <int NS>
__device__ __forceinline__ void detect_strip(const FrameParams& fp, const OctParams& o, int cs, int rs, int s0, int lane) {
    constexpr int ND = NS + 2, NG = NS + 3;
    const int x = cs * 64 + 2 * lane;  // x + 1 < pitch always: the pitch is a multiple of 128 floats
    const int xh = lane == 0 ? max(cs * 64 - 1, 0) : min(cs * 64 + 64, o.w - 1);
    const bool halo_lane = (lane == 0) || (lane == 31);
    const int ys = kBorder + rs * kDetRows;
    const int ye = min(o.h - kBorder, ys + kDetRows);
    const bool ok_a = x >= kBorder && x < o.w - kBorder;
    const bool ok_b = x + 1 >= kBorder && x + 1 < o.w - kBorder;
    const float thr = fp.extrema_thr;
    const size_t ls = (size_t)o.pitch * o.h;
    const float* col = o.gauss + ls * s0;

    float pmx[ND][2], pmn[ND][2];  // horizontal 3-max / 3-min of the previous row
    float qmx[ND][2], qmn[ND][2];  // ... of the previous two rows combined
    float cen[NS][2];              // centre-row DoG values of the NS middle layers
    float nmx[ND][2], nmn[ND][2], ncen[NS][2];

    // horizontal pass of one raw row -> nmx/nmn/ncen
    auto horizontal = [&](const RawRow<NG>& r) {
#pragma unroll
        for (int l = 0; l < ND; l++) {
            const float da = __fsub_rn(r.g[l + 1].x, r.g[l].x), db = __fsub_rn(r.g[l + 1].y, r.g[l].y);
            const float dh = __fsub_rn(r.gh[l + 1], r.gh[l]);
            float left = __shfl_up_sync(0xffffffffu, db, 1);
            float right = __shfl_down_sync(0xffffffffu, da, 1);
            if (lane == 0) left = dh;
            if (lane == 31) right = dh;
            const float hi = fmaxf(da, db), lo = fminf(da, db);
            nmx[l][0] = fmaxf(left, hi); nmx[l][1] = fmaxf(hi, right);
            nmn[l][0] = fminf(left, lo); nmn[l][1] = fminf(lo, right);
            if (l >= 1 && l <= NS) { ncen[l - 1][0] = da; ncen[l - 1][1] = db; }
        }
    };

    RawRow<NG> raw;
    load_raw<NG>(col + (size_t)(ys - 1) * o.pitch, ls, x, xh, halo_lane, raw);
    horizontal(raw);
#pragma unroll
    for (int l = 0; l < ND; l++)
#pragma unroll
        for (int k = 0; k < 2; k++) { pmx[l][k] = nmx[l][k]; pmn[l][k] = nmn[l][k]; }
    load_raw<NG>(col + (size_t)ys * o.pitch, ls, x, xh, halo_lane, raw);
    horizontal(raw);
#pragma unroll
    for (int l = 0; l < ND; l++)
#pragma unroll
        for (int k = 0; k < 2; k++) {
            qmx[l][k] = fmaxf(pmx[l][k], nmx[l][k]); qmn[l][k] = fminf(pmn[l][k], nmn[l][k]);
            pmx[l][k] = nmx[l][k]; pmn[l][k] = nmn[l][k];
        }
#pragma unroll
    for (int s = 0; s < NS; s++) { cen[s][0] = ncen[s][0]; cen[s][1] = ncen[s][1]; }
    load_raw<NG>(col + (size_t)(ys + 1) * o.pitch, ls, x, xh, halo_lane, raw);

    for (int y = ys; y < ye; y++) {
        // `raw` holds row y+1; start fetching row y+2 (the last one is a harmless in-bounds re-read)
        RawRow<NG> nxt;
        load_raw<NG>(col + (size_t)min(y + 2, o.h - 1) * o.pitch, ls, x, xh, halo_lane, nxt);
        horizontal(raw);
        float vmx[ND][2], vmn[ND][2];
#pragma unroll
        for (int l = 0; l < ND; l++)
#pragma unroll
            for (int k = 0; k < 2; k++) {
                vmx[l][k] = fmaxf(qmx[l][k], nmx[l][k]); vmn[l][k] = fminf(qmn[l][k], nmn[l][k]);
                qmx[l][k] = fmaxf(pmx[l][k], nmx[l][k]); qmn[l][k] = fminf(pmn[l][k], nmn[l][k]);
                pmx[l][k] = nmx[l][k]; pmn[l][k] = nmn[l][k];
            }
        unsigned be[NS], bo[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) {
            bool ext[2];
#pragma unroll
            for (int k = 0; k < 2; k++) {
                const float v = cen[s][k];
                const float m3 = fmaxf(fmaxf(vmx[s][k], vmx[s + 1][k]), vmx[s + 2][k]);
                const float n3 = fminf(fminf(vmn[s][k], vmn[s + 1][k]), vmn[s + 2][k]);
                // MatOps.cu:122-143: |v| > thr, and v >= all 27 (v > 0) or v <= all 27 (v < 0).
                ext[k] = (k == 0 ? ok_a : ok_b) && (fabsf(v) > thr) && ((v > 0.f) ? (v >= m3) : ((v < 0.f) && (v <= n3)));
            }
            be[s] = __ballot_sync(0xffffffffu, ext[0]);
            bo[s] = __ballot_sync(0xffffffffu, ext[1]);
            cen[s][0] = ncen[s][0]; cen[s][1] = ncen[s][1];
        }
        if (lane < NS) {  // lane s stores scale s0 + s
            unsigned e = be[0], od = bo[0];
#pragma unroll
            for (int s = 1; s < NS; s++)
                if (lane == s) { e = be[s]; od = bo[s]; }
            const int row = (s0 + lane) * o.h + y;
            *reinterpret_cast<uint2*>(o.mask + (size_t)row * o.words + 2 * cs) = make_uint2(e, od);
            if (e | od) atomicAdd(o.rowcount + row, __popc(e) + __popc(od));
        }
        raw = nxt;
    }
}

// grid.x: warp tasks (strip x row block) of the octave range, grid.y: groups of up to three scales -- scale group g
// covers scales [3g, min(nScales, 3g + 3)) and reads the Gaussian layers 3g .. 3g + ns + 2 (numOctaveLayers = 3: one
// group, every layer read once).
__global__ void __launch_bounds__(kDetThreads, 4)
detect_kernel(const __grid_constant__ FrameParams fp) {
    const int lane = threadIdx.x & 31;
    const int task = fp.oct[fp.o_begin].det_task0 + blockIdx.x * (kDetThreads / 32) + (threadIdx.x >> 5);
    if (task >= (fp.o_end < fp.nOct ? fp.oct[fp.o_end].det_task0 : fp.det_tasks)) return;
    int oi = fp.o_begin;
    while (oi + 1 < fp.o_end && task >= fp.oct[oi + 1].det_task0) oi++;
    const OctParams& o = fp.oct[oi];
    const int t = task - o.det_task0;
    const int cs = t % o.det_cols, rs = t / o.det_cols;
    const int s0 = 3 * blockIdx.y;
    switch (min(3, fp.nScales - s0)) {
        case 3: detect_strip<3>(fp, o, cs, rs, s0, lane); break;
        case 2: detect_strip<2>(fp, o, cs, rs, s0, lane); break;
        case 1: detect_strip<1>(fp, o, cs, rs, s0, lane); break;
        default: break;
    }
}

// CTA per 8 consecutive (scale,row) entries of one octave, warp per row: append the row's extrema in x order.
// The row's offset into the candidate list is the sum of the counts of all earlier rows; every CTA adds
// those up itself (a few thousand L2-resident integers) instead of waiting for a separate scan kernel.
__global__ void __launch_bounds__(256)
scatter_kernel(const __grid_constant__ FrameParams fp) {
    __shared__ int s_part[8], s_cnt[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cta = fp.oct[fp.o_begin].scat_cta0 + blockIdx.x;
    int oi = fp.o_begin;
    while (oi + 1 < fp.o_end && cta >= fp.oct[oi + 1].scat_cta0) oi++;
    const OctParams& o = fp.oct[oi];
    const int nrows = fp.nScales * o.h;
    const int row_first = (cta - o.scat_cta0) * 8;
    const int row = row_first + warp;
    int part = 0;
    for (int i = threadIdx.x; i < row_first; i += 256) part += o.rowcount[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
    const int cnt = row < nrows ? o.rowcount[row] : 0;
    if (lane == 0) { s_part[warp] = part; s_cnt[warp] = cnt; }
    __syncthreads();
    int base = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) base += s_part[w] + (w < warp ? s_cnt[w] : 0);
    if (row == nrows - 1 && lane == 0) {
        o.counts[0] = base + cnt;               // P
        o.counts[3] = min(base + cnt, o.cap);   // candidates the refinement looks at
    }
    if (cnt == 0) return;
    const int s = row / o.h, y = row - s * o.h;
    const uint2* m = reinterpret_cast<const uint2*>(o.mask + (size_t)row * o.words);
    const int strips = o.words >> 1;
    for (int w0 = 0; w0 < strips; w0 += 32) {
        const int wi = w0 + lane;  // 64-pixel strip: .x = even pixels, .y = odd pixels (detect_kernel)
        uint2 bits = wi < strips ? m[wi] : make_uint2(0u, 0u);
        const int c = __popc(bits.x) + __popc(bits.y);
        int incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int u = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += u;
        }
        int idx = base + incl - c;
        while (bits.x | bits.y) {
            const int be = bits.x ? __ffs(bits.x) - 1 : 32, bo = bits.y ? __ffs(bits.y) - 1 : 32;
            int xx;
            if (be <= bo) { xx = wi * 64 + 2 * be; bits.x &= bits.x - 1; }
            else { xx = wi * 64 + 2 * bo + 1; bits.y &= bits.y - 1; }
            if (idx + 1 < o.cap) o.cand[idx] = make_float3((float)xx, (float)y, (float)(s + 1));
            idx++;
        }
        base += __shfl_sync(0xffffffffu, incl, 31);
    }
}

// ---------------------------------------------------------------- refinement
struct DogView {
    const float* g;
    int pitch;
    size_t ls;
    __device__ __forceinline__ float operator()(int l, int r, int c) const {
        const float* p = g + ls * l + (size_t)r * pitch + c;
        return __fsub_rn(__ldg(p + ls), __ldg(p));
    }
};

// solve_3x3_system (SiftOps.cu:6-61): partial pivoting; every `x -= t*y` is one FFMA, divisions
// are IEEE (the reference is built without fast-math).
__device__ __forceinline__ void solve3(const float* A, const float* b, float* x) {
    float a[3][3], bb[3], t;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        bb[i] = b[i];
#pragma unroll
        for (int j = 0; j < 3; j++) a[i][j] = A[i * 3 + j];
    }
#pragma unroll
    for (int k = 0; k < 2; k++) {
        int mi = k;
        float mv = fabsf(a[k][k]);
#pragma unroll
        for (int i = k + 1; i < 3; i++)
            if (fabsf(a[i][k]) > mv) { mv = fabsf(a[i][k]); mi = i; }
        if (mi != k) {
#pragma unroll
            for (int i = k + 1; i < 3; i++)
                if (i == mi) {
#pragma unroll
                    for (int j = k; j < 3; j++) { t = a[k][j]; a[k][j] = a[i][j]; a[i][j] = t; }
                    t = bb[k]; bb[k] = bb[i]; bb[i] = t;
                }
        }
#pragma unroll
        for (int i = k + 1; i < 3; i++) {
            t = __fdiv_rn(a[i][k], a[k][k]);
            bb[i] = __fmaf_rn(-t, bb[k], bb[i]);
#pragma unroll
            for (int j = k; j < 3; j++) a[i][j] = __fmaf_rn(-t, a[k][j], a[i][j]);
        }
    }
    x[2] = __fdiv_rn(bb[2], a[2][2]);
    x[1] = __fdiv_rn(__fmaf_rn(-a[1][2], x[2], bb[1]), a[1][1]);
    x[0] = __fdiv_rn(__fmaf_rn(-a[0][2], x[2], __fmaf_rn(-a[0][1], x[1], bb[0])), a[0][0]);
}

__global__ void __launch_bounds__(128)
refine_kernel(const __grid_constant__ FrameParams fp) {
    const int oi = fp.o_begin + blockIdx.y;
    const OctParams& o = fp.oct[oi];
    const int idx = blockIdx.x * 128 + threadIdx.x;
    if (idx >= o.counts[3]) return;
    const float3 kp = o.cand[idx];
    int c = (int)kp.x, r = (int)kp.y, l = (int)kp.z;
    const float sx = (float)o.w, sy = (float)o.h;
    constexpr float img_scale = 1.f / 255.f;
    constexpr float deriv_scale = img_scale * 0.5f, second = img_scale, cross = img_scale * 0.25f;
    const DogView D{o.gauss, o.pitch, (size_t)o.pitch * o.h};
    float A[9], b[3], x[3] = {0.f, 0.f, 0.f};
    int it = 0;
    int accept = 1;
    // A stale slot (only reachable when an octave overflows its cap, appendix A.1) can hold
    // anything; the reference would read outside its DoG block there. Reject instead.
    const int nS = fp.nScales;
    if (l < 1 || l > nS || r < 1 || r >= o.h - 1 || c < 1 || c >= o.w - 1) { accept = 0; it = 5; }
    for (; it < 5; it++) {
        const float c_c = D(l, r, c);
        const float c_xp = D(l, r, c + 1), c_xm = D(l, r, c - 1);
        const float c_yp = D(l, r + 1, c), c_ym = D(l, r - 1, c);
        const float n_c = D(l + 1, r, c), p_c = D(l - 1, r, c);
        b[0] = __fmul_rn(deriv_scale, __fsub_rn(c_xp, c_xm));
        b[1] = __fmul_rn(deriv_scale, __fsub_rn(c_yp, c_ym));
        b[2] = __fmul_rn(deriv_scale, __fsub_rn(n_c, p_c));
        const float v2 = __fadd_rn(c_c, c_c);
        A[0] = __fmul_rn(__fsub_rn(__fadd_rn(c_xp, c_xm), v2), second);
        A[4] = __fmul_rn(__fsub_rn(__fadd_rn(c_yp, c_ym), v2), second);
        A[8] = __fmul_rn(__fsub_rn(__fadd_rn(n_c, p_c), v2), second);
        const float dxy = __fmul_rn(__fadd_rn(__fsub_rn(__fsub_rn(D(l, r + 1, c + 1), D(l, r + 1, c - 1)), D(l, r - 1, c + 1)), D(l, r - 1, c - 1)), cross);
        const float dxs = __fmul_rn(__fadd_rn(__fsub_rn(__fsub_rn(D(l + 1, r, c + 1), D(l + 1, r, c - 1)), D(l - 1, r, c + 1)), D(l - 1, r, c - 1)), cross);
        const float dys = __fmul_rn(__fadd_rn(__fsub_rn(__fsub_rn(D(l + 1, r + 1, c), D(l + 1, r - 1, c)), D(l - 1, r + 1, c)), D(l - 1, r - 1, c)), cross);
        A[1] = A[3] = dxy;
        A[2] = A[6] = dxs;
        A[5] = A[7] = dys;
        solve3(A, b, x);
        if (fabsf(x[0]) < 0.5f && fabsf(x[1]) < 0.5f && fabsf(x[2]) < 0.5f) break;
        if (fabsf(x[0]) > sx || fabsf(x[1]) > sy || fabsf(x[2]) > 100.f) { accept = 0; break; }
        l = (int)__fsub_rn((float)l, roundf(x[2]));
        r = (int)__fsub_rn((float)r, roundf(x[1]));
        c = (int)__fsub_rn((float)c, roundf(x[0]));
        if (l < 1 || l > nS) { accept = 0; break; }  // SiftOps.cu:154: l > size.z - 2, size.z = n + 2 DoG layers
        if ((float)r < 5.f || (float)r >= sy - 5.f || (float)c < 5.f || (float)c >= sx - 5.f) { accept = 0; break; }
    }
    if (it >= 5) accept = 0;
    if (accept) {
        const float c_c = D(l, r, c);
        const float c_xp = D(l, r, c + 1), c_xm = D(l, r, c - 1);
        const float c_yp = D(l, r + 1, c), c_ym = D(l, r - 1, c);
        const float d0 = __fmul_rn(deriv_scale, __fsub_rn(c_xp, c_xm));
        const float d1 = __fmul_rn(deriv_scale, __fsub_rn(c_yp, c_ym));
        const float d2 = __fmul_rn(deriv_scale, __fsub_rn(D(l + 1, r, c), D(l - 1, r, c)));
        // reference SASS: t = fma(d2,x2, fma(d0,x0, d1*x1)); contrast = fma(v, 1/255, t*-0.5)
        const float t = __fmaf_rn(d2, x[2], __fmaf_rn(d0, x[0], __fmul_rn(d1, x[1])));
        const float contrast = __fmaf_rn(c_c, img_scale, __fmul_rn(t, -0.5f));
        if (__fmul_rn(fabsf(contrast), (float)nS) < fp.contrast_thr) accept = 0;
        if (accept) {
            const float v2 = __fadd_rn(c_c, c_c);
            const float dxx = __fmul_rn(__fsub_rn(__fadd_rn(c_xp, c_xm), v2), second);
            const float dyy = __fmul_rn(__fsub_rn(__fadd_rn(c_yp, c_ym), v2), second);
            const float dxy = __fmul_rn(__fadd_rn(__fsub_rn(__fsub_rn(D(l, r + 1, c + 1), D(l, r + 1, c - 1)), D(l, r - 1, c + 1)), D(l, r - 1, c - 1)), cross);
            const float tr = __fadd_rn(dxx, dyy);
            const float det = __fmaf_rn(dxx, dyy, -__fmul_rn(dxy, dxy));
            const float e = fp.edge_thr;
            const float e1 = __fadd_rn(e, 1.f);
            if (det <= 0.f || __fmul_rn(__fmul_rn(tr, tr), e) >= __fmul_rn(__fmul_rn(e1, e1), det)) accept = 0;
        }
        if (accept) {
            // SiftOps.cu:201-207 (quirk Q10: integer position, sub-pixel offset not added)
            // opt_upscale (working upscale, opt-in): octave oi of the doubled frame is octave oi - 1 of the input:
            // positions halved (as the reference does, SiftOps.cu:200-202), and -- the fix -- size and packed octave too
            const float inv_mod = fp.opt_upscale ? 0.5f : 1.f;
            o.cand[idx] = make_float3(__fmul_rn((float)(c * (1 << oi)), inv_mod), __fmul_rn((float)(r * (1 << oi)), inv_mod), (float)l);
            float4 f = o.candF[idx];
            const int octave_field = fp.opt_upscale ? ((oi - 1) & 255) : oi;
            f.x = (float)(octave_field + (l << 8) + ((int)__fmul_rn(roundf(__fadd_rn(x[2], 0.5f)), 255.f) << 16));
            const float pw = powf(2.f, __fdiv_rn(__fsub_rn((float)l, x[2]), (float)nS));
            const float sp = __fmul_rn(__fmul_rn(fp.sigma, pw), (float)(1 << oi));
            f.y = __fmul_rn(__fadd_rn(sp, sp), inv_mod);
            f.z = fabsf(contrast);
            o.candF[idx] = f;
        }
    }
    o.flag[idx] = accept;
}

// CTA (blockIdx.x, octave): stable compaction of the accepted candidates of its kScanBlock-element segment into
// coll/collF. The element whose prefix equals the total is not written (quirk Q5, SiftOps.cu:228).
__global__ void __launch_bounds__(kScanBlock)
compact_refined_kernel(const __grid_constant__ FrameParams fp) {
    const OctParams& o = fp.oct[fp.o_begin + blockIdx.y];
    const int n = o.counts[3];
    if (blockIdx.x > 0 && (int)blockIdx.x * kScanBlock >= n) return;
    const int total = segment_scan(
        n, blockIdx.x, [&](int i) { return o.flag[i]; },
        [&](int i, int v, int excl, int tot) {
            if (v && excl + 1 < tot) {
                o.coll[excl] = o.cand[i];
                o.collF[excl] = o.candF[i];
            }
        });
    if (blockIdx.x == 0 && threadIdx.x == 0) o.counts[1] = total;  // M
}

}  // namespace

void launch_detect(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s) {
    const int warps = kDetThreads / 32;
    const int t0 = fp.oct[o_begin].det_task0, t1 = o_end < fp.nOct ? fp.oct[o_end].det_task0 : fp.det_tasks;
    dim3 grid((t1 - t0 + warps - 1) / warps, (fp.nScales + 2) / 3);
    if (t1 > t0) detect_kernel<<<grid, kDetThreads, 0, s>>>(octave_range(fp, o_begin, o_end));
}
void launch_scatter(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s) {
    const OctParams& last = fp.oct[o_end - 1];
    const int c0 = fp.oct[o_begin].scat_cta0, c1 = last.scat_cta0 + (fp.nScales * last.h + 7) / 8;
    scatter_kernel<<<c1 - c0, 256, 0, s>>>(octave_range(fp, o_begin, o_end));
}
void launch_refine(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s) {
    int maxcap = 1;
    for (int i = o_begin; i < o_end; i++) maxcap = fp.oct[i].cap > maxcap ? fp.oct[i].cap : maxcap;
    dim3 grid((maxcap + 127) / 128, o_end - o_begin);
    refine_kernel<<<grid, 128, 0, s>>>(octave_range(fp, o_begin, o_end));
}
void launch_compact_refined(const FrameParams& fp, int o_begin, int o_end, cudaStream_t s) {
    int maxcap = 1;
    for (int i = o_begin; i < o_end; i++) maxcap = fp.oct[i].cap > maxcap ? fp.oct[i].cap : maxcap;
    dim3 grid((maxcap + kScanBlock - 1) / kScanBlock, o_end - o_begin);
    compact_refined_kernel<<<grid, kScanBlock, 0, s>>>(octave_range(fp, o_begin, o_end));
}

}  // namespace b200sift
