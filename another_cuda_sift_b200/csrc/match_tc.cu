// Tensor-core brute-force matcher for sm_100a: tcgen05.mma + TMEM candidate filter, exact re-rank.
//
// The reference (sift_func/Match.cu:35-177) scores every (query, train) pair with rounded-half arithmetic
//     score = sum_j half( d_j*d_j ),  d_j = half( half(a_j - b_j) * 0.25 )            (~ |a-b|^2 / 16)
// keeps, per query, the top-2 of each of 32 contiguous column chunks (strict '<', lowest index wins) and
// merges the 32 results with a shuffle tree (quirk Q15). Only the top-2 of every (query, chunk) can
// influence the output, so:
//
//  1. match_filter_kernel: D = |a|^2 + |b|^2 - 2 a.b^T on the 5th-gen tensor cores (tcgen05.mma
//     kind::f16, 128 x 128 x 16 per instruction, fp32 accumulators in TMEM). Inputs are exact halves, so
//     D is the true squared distance up to fp32 accumulation error. Every half rounding of the reference
//     recipe is a relative error <= 2^-11 on a non-negative term, hence |16*score - D| <= 1.5e-3 * D
//     (+ tiny absolute slack). Pass 1 finds the 2nd smallest D of the chunk, pass 2 emits every column
//     with D <= D2nd * 1.004 + eps: a superset of the exact top-2 including ties. Typically 2-3 columns
//     per (query, chunk) survive.
//  2. match_rerank_kernel: warp per query re-scores the survivors with the reference's exact arithmetic
//     (lane i owns half2 i and i+32, same fp32 tree), in ascending column order with the same strict-'<'
//     updates, then replays the reference's 32-lane merge and the double-precision ratio test.
//     A (query, chunk) whose survivor list overflows is re-scored column by column (rare, still exact).
//
// Roofline: tensor-bound in the limit (2*nq*nt*128 flop); at 2k x 2k it is launch/latency-bound.
#include <algorithm>
#include <cuda.h>
#include "common.cuh"

namespace b200sift {

namespace {

constexpr int kTileM = 128;   // queries per CTA  (= TMEM lanes)
constexpr int kCand = 7;      // survivors kept per (query, chunk); record = {count, c0..c6}
constexpr int kFilterThreads = 128;
constexpr uint32_t kTileBytes = kTileM * 256;  // 128 rows x 128 halves
constexpr uint32_t kKcStride = kTileM * 16;    // bytes between consecutive 8-column K chunks

struct __align__(16) CandRec {
    int n;            // 0..kCand, or kCand+1 = overflow (re-score the whole chunk)
    int c[kCand];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, no-swizzle UMMA shared-memory descriptor. Tile layout in smem: [k_chunk(16)][row(128)][8 halves]:
// a core matrix is 8 rows x 16 B = 128 contiguous bytes; next 8-row group +128 B (SBO); next 8-column
// K chunk +128*16 B (LBO).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
    return d;                // base_offset 0, lbo_mode 0, layout_type 0 = SWIZZLE_NONE
}

// kind::f16 instruction descriptor: D=f32, A=B=f16, both K-major, M=128, N=n.
__host__ __device__ constexpr uint32_t make_idesc(int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}

// One lane of a converged warp. ptxas knows that a region guarded by elect.sync runs a single thread and keeps
// the operands of the tcgen05 / TMA instructions inside it in uniform registers; guarded by `lane == 0` every such
// instruction is wrapped in an ELECT + R2UR.BROADCAST uniformisation loop (13 issue slots per MMA, measured).
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spins = 0; !done; spins++) {
        if (spins > (1u << 28)) __trap();  // a lost commit must fail loudly, never hang the GPU
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    }
}

// TR rows x 256 B from global (row-major) into the UMMA layout [k_chunk][row(TR)][16 B]. Rows in
// [rows, round_up(rows, 32)) are zero-filled (their distances become +inf through s_nb and are scanned);
// rows beyond that are left as they are: their products land in TMEM columns nobody reads.
template <int TR>
__device__ __forceinline__ void load_tile(const __half* __restrict__ g, int rows, uint8_t* s, int warp, int lane) {
    const int r8 = lane & 7, k4 = lane >> 3;
    const int rows32 = (rows + 31) & ~31;
    constexpr int kBatch = 16;  // 16-byte loads in flight per thread
#pragma unroll 1
    for (int u0 = warp; u0 < TR / 2; u0 += kBatch * (kFilterThreads / 32)) {
        uint4 v[kBatch];
#pragma unroll
        for (int i = 0; i < kBatch; i++) {
            const int u = u0 + i * (kFilterThreads / 32);
            const int row = (u >> 2) * 8 + r8, kc = (u & 3) * 4 + k4;
            v[i] = make_uint4(0u, 0u, 0u, 0u);
            if (row < rows) v[i] = __ldg(reinterpret_cast<const uint4*>(g + (size_t)row * 128) + kc);
        }
#pragma unroll
        for (int i = 0; i < kBatch; i++) {
            const int u = u0 + i * (kFilterThreads / 32);
            const int row = (u >> 2) * 8 + r8, kc = (u & 3) * 4 + k4;
            if (row < rows32) *reinterpret_cast<uint4*>(s + kc * (TR * 16) + row * 16) = v[i];
        }
    }
}

template <int TR>
__device__ __forceinline__ float row_norm2(const uint8_t* s, int row) {
    float acc = 0.f;
#pragma unroll
    for (int kc = 0; kc < 16; kc++) {
        const uint4 v = *reinterpret_cast<const uint4*>(s + kc * (TR * 16) + row * 16);
        const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float2 f = __half22float2(h[i]);
            acc = __fmaf_rn(f.x, f.x, acc);
            acc = __fmaf_rn(f.y, f.y, acc);
        }
    }
    return acc;
}

// TN = train rows per MMA tile (= TMEM columns). When the whole chunk fits one tile (per <= TN) the
// distances stay resident in TMEM and pass 2 only re-reads them (no reload, no second MMA).
template <int TN, bool DBG>
__global__ void __launch_bounds__(kFilterThreads)
match_filter_kernel(const __half* __restrict__ des, int nq, const __half* __restrict__ src, int nt, int per,
                    CandRec* __restrict__ rec, float* __restrict__ dbg, const int* __restrict__ cnt_q,
                    const int* __restrict__ cnt_t) {
    if (cnt_q) {  // sizes known only on the device (consecutive-frame matching without a host round trip)
        nt = *cnt_t;
        nq = *cnt_q;
        per = (int)ceilf((float)nt / 32.f);
    }
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sQ = smem;
    uint8_t* sB = smem + kTileBytes;
    float* s_nb = reinterpret_cast<float*>(smem + kTileBytes + TN * 256);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_nb + TN);
    constexpr uint32_t kIdesc = make_idesc(TN);
    constexpr uint32_t kKcStrideB = TN * 16;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 1);
    int* s_nbmax = reinterpret_cast<int*>(s_tmem + 1);
    int* s_cand = reinterpret_cast<int*>(s_tmem + 4);  // [kCand][128]: survivor columns of thread tid

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q0 = blockIdx.x * kTileM;
    const int chunk = blockIdx.y;
    const int t_begin = chunk * per;
    const int t_end = min(nt, t_begin + per);
    const int q = q0 + tid;
    if (q0 >= nq) return;    // grid was sized for the largest possible query set
    if (t_begin >= t_end) {  // empty chunk (uniform for the CTA): nothing to score
        if (q < nq) rec[(size_t)q * 32 + chunk].n = 0;
        return;
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        *s_nbmax = 0;
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    load_tile<kTileM>(des + (size_t)q0 * 128, min(kTileM, nq - q0), sQ, warp, lane);
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = *s_tmem;
    const float na = (q < nq) ? row_norm2<kTileM>(sQ, tid) : 0.f;
    const bool resident = per <= TN;

    const uint32_t bar = smem_u32(s_bar);
    const uint32_t aQ = smem_u32(sQ), aB = smem_u32(sB);
    uint32_t phase = 0;
    float a1 = INFINITY, a2 = INFINITY, thr = 0.f;
    int cn = 0;

    for (int pass = 0; pass < 2; pass++) {
        for (int t0 = t_begin; t0 < t_end; t0 += TN) {
            const int rows = min(TN, t_end - t0);
            if (!(resident && pass == 1)) {
            load_tile<TN>(src + (size_t)t0 * 128, rows, sB, warp, lane);
            __syncthreads();
            for (int r = tid; r < ((rows + 31) & ~31); r += kFilterThreads) {
                const float nb = r < rows ? row_norm2<TN>(sB, r) : INFINITY;  // padded columns can never win
                s_nb[r] = nb;
                if (r < rows) atomicMax(s_nbmax, __float_as_int(nb));         // norms are >= 0: int order == float order
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // st.shared -> visible to the tensor core
            __syncthreads();
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;");
#pragma unroll
                for (int k = 0; k < 8; k++) {  // K = 128 = 8 x 16; one instruction consumes 2 K chunks
                    const uint64_t da = umma_desc(aQ + k * 2 * kKcStride, kKcStride, 128);
                    const uint64_t db = umma_desc(aB + k * 2 * kKcStrideB, kKcStrideB, 128);
                    const uint32_t acc = k > 0 ? 1u : 0u;
                    asm volatile(
                        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                        ::"r"(tmem), "l"(da), "l"(db), "r"(kIdesc), "r"(acc) : "memory");
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
            }
            mbar_wait(bar, phase);
            phase ^= 1;
            asm volatile("tcgen05.fence::after_thread_sync;");
            }
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
            for (int cb = 0; cb < TN / 32; cb++) {
                if (cb * 32 >= rows) break;  // warp-uniform
                uint32_t v[32];
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr + cb * 32));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (pass == 0) {
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        const float d = __fadd_rn(na, __fmaf_rn(-2.f, __uint_as_float(v[j]), s_nb[cb * 32 + j]));
                        a2 = fminf(a2, fmaxf(a1, d));  // branch-free running top-2
                        a1 = fminf(a1, d);
                        if (DBG) { if (q < nq && cb * 32 + j < rows) dbg[(size_t)q * nt + t0 + cb * 32 + j] = d; }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        const float d = __fadd_rn(na, __fmaf_rn(-2.f, __uint_as_float(v[j]), s_nb[cb * 32 + j]));
                        if (d <= thr) {  // survivors are rare: one predicated shared-memory store each
                            if (cn < kCand) s_cand[cn * kFilterThreads + tid] = t0 + cb * 32 + j;
                            cn++;
                        }
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;");
            __syncthreads();  // sB, s_nb and the TMEM tile are free for the next tile
        }
        // survivors: D <= D2nd * (1+d)/(1-d) with d = 1.5e-3 (half roundings), + fp32 accumulation slack
        thr = __fmaf_rn(fmaxf(a2, 0.f), 1.004f, __fmaf_rn(4e-5f, __fadd_rn(na, __int_as_float(*s_nbmax)), 1e-3f));
    }
    if (q < nq) {
        int4* dst = reinterpret_cast<int4*>(rec + (size_t)q * 32 + chunk);
        int c[kCand];
#pragma unroll
        for (int i = 0; i < kCand; i++) c[i] = i < cn ? s_cand[i * kFilterThreads + tid] : -1;
        dst[0] = make_int4(min(cn, kCand + 1), c[0], c[1], c[2]);
        dst[1] = make_int4(c[3], c[4], c[5], c[6]);
    }
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TN));
    }
}

// ---------------------------------------------------------------- resident-chunk filter (per <= 256)
// Squared norms of every row (fp32), once per matcher call instead of once per CTA.
__global__ void __launch_bounds__(256)
match_norms_kernel(const __half* __restrict__ des, int nq, const __half* __restrict__ src, int nt, float* __restrict__ nrm_q,
                   float* __restrict__ nrm_t, const int* __restrict__ cnt_q, const int* __restrict__ cnt_t) {
    if (cnt_q) { nt = *cnt_t; nq = *cnt_q; }
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= nq + nt) return;
    const __half* p = row < nq ? des + (size_t)row * 128 : src + (size_t)(row - nq) * 128;
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p) + lane);
    const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
    const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
    float acc = __fmaf_rn(f0.x, f0.x, __fmaf_rn(f0.y, f0.y, __fmaf_rn(f1.x, f1.x, __fmul_rn(f1.y, f1.y))));
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if (lane == 0) (row < nq ? nrm_q[row] : nrm_t[row - nq]) = acc;
}

// Streaming filter only: the ninth K step of its MMA. Per train row one 16-byte K chunk {h1, h2, h3, 0, 0, 0, 0, 0}
// with |b|^2 == 2^15*h1 + 2^4*h2 + 2^-6*h3 EXACTLY (three halves cover the 24-bit fp32 mantissa; every residual is
// an exact fp32 subtraction), so that the MMA itself adds |b|^2: the query side carries {2^15, 2^4, 2^-6, 0...} in
// the same chunk. Representable while |b|^2 <= 65504 * 2^15 (~2.1e9); the filter falls back to the exact path
// above that. Laid out per (chunk, 128-row tile) exactly as the tile sits in shared memory -- [2 K chunks][128
// rows][16 B], the second K chunk zero -- so one 4 KB bulk copy brings it in; rows past the chunk (or past the
// train set) carry {+inf, 0...}: their distance must lose.
constexpr float kAugS1 = 32768.f, kAugS2 = 16.f, kAugS3 = 1.f / 64.f;
__global__ void __launch_bounds__(256)
match_aug_kernel(const float* __restrict__ nrm_t, int nt, int per, int tiles, uint4* __restrict__ aug) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= 32 * tiles * 128) return;
    const int c = idx / (tiles * 128), rr = idx - c * (tiles * 128);
    const int t = c * per + rr;
    uint4 o = make_uint4(0x7c00u, 0u, 0u, 0u);  // {+inf, 0, ...}
    if (rr < per && t < nt) {
        const float nb = __ldg(nrm_t + t);
        const __half h1 = __float2half_rn(__fmul_rn(nb, 1.f / kAugS1));
        const float r1 = __fsub_rn(nb, __fmul_rn(__half2float(h1), kAugS1));
        const __half h2 = __float2half_rn(__fmul_rn(r1, 1.f / kAugS2));
        const float r2 = __fsub_rn(r1, __fmul_rn(__half2float(h2), kAugS2));
        const __half h3 = __float2half_rn(__fmul_rn(r2, 1.f / kAugS3));
        o.x = (uint32_t)__half_as_ushort(h1) | ((uint32_t)__half_as_ushort(h2) << 16);
        o.y = (uint32_t)__half_as_ushort(h3);
    }
    uint4* tile = aug + ((size_t)c * tiles + (rr >> 7)) * 256;
    tile[rr & 127] = o;
    tile[128 + (rr & 127)] = make_uint4(0u, 0u, 0u, 0u);
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int bytes = valid ? 16 : 0;  // src-size 0: the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}

constexpr int kResThreads = 256;  // two groups of four warps, 128 queries each, sharing every train tile
constexpr int kResQ = 256;        // queries per CTA

// One CTA = 256 queries x `cpc` consecutive column chunks of the reference (chunk c = columns [c*per, (c+1)*per)).
// Every chunk fits one MMA tile (TR = per rounded up to 32 <= 256 rows), so its distances stay resident in
// TMEM (two 128 x TR accumulators, one per query half): pass 1 (second smallest) and pass 2 (survivors) both
// read TMEM, one MMA pair per chunk. The query tiles are loaded once per CTA, pre-multiplied by -2 (exact in
// fp16), so the accumulator is -2 a.b and a distance is one FADD with the precomputed |b|^2; |a|^2 is
// constant per thread and only enters the threshold. The next chunk's train tile streams in with cp.async
// while the current one is scanned.
#define TMEM_LD32(v, addr)                                                                                     \
    asm volatile(                                                                                              \
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                              \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                              \
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"              \
        : "=r"((v)[0]), "=r"((v)[1]), "=r"((v)[2]), "=r"((v)[3]), "=r"((v)[4]), "=r"((v)[5]), "=r"((v)[6]), "=r"((v)[7]),          \
          "=r"((v)[8]), "=r"((v)[9]), "=r"((v)[10]), "=r"((v)[11]), "=r"((v)[12]), "=r"((v)[13]), "=r"((v)[14]), "=r"((v)[15]),    \
          "=r"((v)[16]), "=r"((v)[17]), "=r"((v)[18]), "=r"((v)[19]), "=r"((v)[20]), "=r"((v)[21]), "=r"((v)[22]), "=r"((v)[23]),  \
          "=r"((v)[24]), "=r"((v)[25]), "=r"((v)[26]), "=r"((v)[27]), "=r"((v)[28]), "=r"((v)[29]), "=r"((v)[30]), "=r"((v)[31])   \
        : "r"(addr))

// NB > 0: the chunk tile has exactly NB 32-column blocks (TR = 32*NB <= 160) and a thread keeps all its
// distances in registers between the two passes (one TMEM read). NB == 0: any TR, TMEM is read twice.
template <bool DBG, int NB>
__global__ void __launch_bounds__(kResThreads)
match_filter_resident_kernel(const __half* __restrict__ des, int nq, const __half* __restrict__ src, int nt, int per, int TR, int cpc,
                             CandRec* __restrict__ rec, float* __restrict__ dbg, const float* __restrict__ nrm_q,
                             const float* __restrict__ nrm_t, const int* __restrict__ cnt_q, const int* __restrict__ cnt_t) {
    if (cnt_q) {
        nt = *cnt_t;
        nq = *cnt_q;
        per = (int)ceilf((float)nt / 32.f);
    }
    const int TRP = TR;  // TR is a multiple of 32: whole 32-column TMEM loads, every scanned column written by the MMA
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sQ = smem;                                     // 2 x [16][128][16 B]
    uint8_t* sB = smem + 2 * kTileBytes;                    // 2 x [16][TR][16 B]
    const uint32_t b_bytes = (uint32_t)TR * 256;
    float* s_nb = reinterpret_cast<float*>(sB + 2 * b_bytes);  // 3 x [TRP]: |b|^2 of the chunk being scanned, the one in the MMA, the one in flight
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_nb + 3 * TRP);
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 1);
    int* s_nbmax = reinterpret_cast<int*>(s_tmem + 1);      // largest |b|^2 over this CTA's chunks (threshold slack)
    unsigned char* s_cand = reinterpret_cast<unsigned char*>(s_tmem + 4);  // [kCand][256]: survivor columns (chunk-relative) of thread tid

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int half_id = warp >> 2;  // which 128-query tile / accumulator this warp works on
    const int q0 = blockIdx.x * kResQ;
    const int q = q0 + tid;
    if (q0 >= nq) return;  // grid was sized for the largest possible query set
    const int c_begin = blockIdx.y * cpc, c_end = min(32, c_begin + cpc);
    const uint32_t idesc = make_idesc(TR);
    const uint32_t kc_stride_b = (uint32_t)TR * 16;
    const bool two_tiles = q0 + kTileM < nq;  // the second query tile holds at least one row

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        s_nbmax[0] = 0;
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    {   // norms are >= 0: int order == float order
        float mx = 0.f;
        for (int t = c_begin * per + tid; t < min(nt, c_end * per); t += kResThreads) mx = fmaxf(mx, __ldg(nrm_t + t));
        atomicMax(s_nbmax, __float_as_int(mx));
    }
    // train tile of chunk c -> buffer (c - c_begin) & 1 (cp.async), its norms -> s_nb[(c - c_begin) % 3]
    // (+inf for the padding columns)
    auto fetch = [&](int c) {
        const int t_begin = c * per;
        const int rows = max(0, min(nt, t_begin + per) - t_begin);
        const uint32_t base = smem_u32(sB) + (uint32_t)((c - c_begin) & 1) * b_bytes;
        const int r8 = lane & 7, k4 = lane >> 3;
        for (int u = warp; u < TR / 2; u += kResThreads / 32) {  // 16-byte pieces: 8 rows x 4 K chunks per warp step
            const int row = (u >> 2) * 8 + r8, kc = (u & 3) * 4 + k4;
            const bool valid = row < rows;
            cp_async16(base + kc * kc_stride_b + row * 16, src + (valid ? (size_t)(t_begin + row) * 128 + kc * 8 : 0), valid);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        float* nbw = s_nb + ((c - c_begin) % 3) * TRP;
        for (int r = tid; r < TRP; r += kResThreads) nbw[r] = r < rows ? __ldg(nrm_t + t_begin + r) : INFINITY;
    };
    fetch(c_begin);
    {   // query tiles, scaled by -2
        const int rows = min(kResQ, nq - q0);
        const int r8 = lane & 7, k4 = lane >> 3;
        const __half2 m2 = __floats2half2_rn(-2.f, -2.f);
        constexpr int kSteps = kResQ / 2 / (kResThreads / 32);  // 16 sixteen-byte pieces per thread, all loads in flight at once
        uint4 v[kSteps];
#pragma unroll
        for (int i = 0; i < kSteps; i++) {
            const int u = warp + i * (kResThreads / 32);
            const int row = (u >> 2) * 8 + r8, kc = (u & 3) * 4 + k4;  // row 0..255
            v[i] = make_uint4(0u, 0u, 0u, 0u);
            if (row < rows) v[i] = __ldg(reinterpret_cast<const uint4*>(des + (size_t)(q0 + row) * 128) + kc);
        }
#pragma unroll
        for (int i = 0; i < kSteps; i++) {
            const int u = warp + i * (kResThreads / 32);
            const int row = (u >> 2) * 8 + r8, kc = (u & 3) * 4 + k4;
            __half2* h = reinterpret_cast<__half2*>(&v[i]);
#pragma unroll
            for (int j = 0; j < 4; j++) h[j] = __hmul2(h[j], m2);
            *reinterpret_cast<uint4*>(sQ + (row >> 7) * kTileBytes + kc * kKcStride + (row & 127) * 16) = v[i];
        }
    }
    const float na = q < nq ? __ldg(nrm_q + q) : 0.f;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = *s_tmem;
    const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)half_id * 256;
    const uint32_t bar = smem_u32(s_bar);
    const uint32_t aQ = smem_u32(sQ);
    uint32_t phase = 0;

    auto chunk_rows = [&](int c) { return max(0, min(nt, c * per + per) - c * per); };
    auto issue_mma = [&](int c) {  // one thread: both query tiles x the train tile of chunk c, then commit
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t aB = smem_u32(sB) + (uint32_t)((c - c_begin) & 1) * b_bytes;
        for (int hq = 0; hq < (two_tiles ? 2 : 1); hq++) {
#pragma unroll
            for (int k = 0; k < 8; k++) {  // K = 128 = 8 x 16; one instruction consumes 2 K chunks
                const uint64_t da = umma_desc(aQ + hq * kTileBytes + k * 2 * kKcStride, kKcStride, 128);
                const uint64_t db = umma_desc(aB + k * 2 * kc_stride_b, kc_stride_b, 128);
                const uint32_t acc = k > 0 ? 1u : 0u;
                asm volatile(
                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                    ::"r"(tmem + hq * 256), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    };
    auto store_rec = [&](int c, int cn) {
        if (q < nq) {
            int4* dst = reinterpret_cast<int4*>(rec + (size_t)q * 32 + c);
            int cc[kCand];
#pragma unroll
            for (int i = 0; i < kCand; i++) cc[i] = i < cn ? c * per + (int)s_cand[i * kResThreads + tid] : -1;
            dst[0] = make_int4(min(cn, kCand + 1), cc[0], cc[1], cc[2]);
            dst[1] = make_int4(cc[3], cc[4], cc[5], cc[6]);
        }
    };
    const float slack = __fmaf_rn(4e-5f, __fadd_rn(na, __int_as_float(s_nbmax[0])), 1e-3f);
    const bool scans = half_id == 0 || two_tiles;
    // top-2 of a value stream with four independent (min, 2nd min) pairs: no serial dependency chain
    float m1[4] = {}, m2[4] = {};
    auto top2 = [&](int k, float dd) {
        m2[k] = fminf(m2[k], fmaxf(m1[k], dd));
        m1[k] = fminf(m1[k], dd);
    };
    auto merged_thr = [&]() {
        // {a1,a2} u {b1,b2}: min = min(a1,b1), 2nd = min(max(a1,b1), min(a2,b2))
        const float x2 = fminf(fmaxf(m1[0], m1[1]), fminf(m2[0], m2[1])), x1 = fminf(m1[0], m1[1]);
        const float y2 = fminf(fmaxf(m1[2], m1[3]), fminf(m2[2], m2[3])), y1 = fminf(m1[2], m1[3]);
        const float a2 = fminf(fmaxf(x1, y1), fminf(x2, y2));
        // survivors: D <= D2nd * (1+d)/(1-d) with d = 1.5e-3 (half roundings), + fp32 accumulation slack;
        // compared in the |b|^2 - 2 a.b domain, i.e. minus |a|^2
        const float d2nd = fmaxf(__fadd_rn(a2, na), 0.f);
        return __fsub_rn(__fmaf_rn(d2nd, 1.004f, slack), na);
    };

    if constexpr (NB > 0) {
        // The distances of a chunk move from TMEM to registers in one go, so the accumulator is free again
        // right away: the MMA of chunk c+1 and the cp.async of chunk c+2 run under the scan of chunk c.
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // cp.async / st.shared -> visible to the tensor core
        __syncthreads();
        if (warp == 0 && chunk_rows(c_begin) > 0) {
            if (elect_one()) issue_mma(c_begin);
            __syncwarp();
        }
        if (c_begin + 1 < c_end) fetch(c_begin + 1);
        for (int c = c_begin; c < c_end; c++) {
            const int rows = chunk_rows(c);
            uint32_t v[NB * 32];
            if (rows > 0) {
                mbar_wait(bar, phase);
                phase ^= 1;
                asm volatile("tcgen05.fence::after_thread_sync;");
                if (scans) {
#pragma unroll
                    for (int cb = 0; cb < NB; cb++) TMEM_LD32(v + cb * 32, taddr + cb * 32);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;");
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();  // accumulator drained by everybody, tile c+1 landed
            if (c + 1 < c_end && warp == 0 && chunk_rows(c + 1) > 0) {
                if (elect_one()) issue_mma(c + 1);
                __syncwarp();
            }
            if (c + 2 < c_end) fetch(c + 2);  // tile buffer of chunk c (its MMA is done), norm slot (c+2) % 3
            int cn = 0;
            if (rows > 0 && scans) {
#pragma unroll
                for (int k = 0; k < 4; k++) { m1[k] = INFINITY; m2[k] = INFINITY; }
                const float4* nb4 = reinterpret_cast<const float4*>(s_nb + ((c - c_begin) % 3) * TRP);
#pragma unroll
                for (int i4 = 0; i4 < NB * 8; i4++) {
                    const float4 nb = nb4[i4];
                    const float nbv[4] = {nb.x, nb.y, nb.z, nb.w};
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const float dd = __fadd_rn(__uint_as_float(v[i4 * 4 + j]), nbv[j]);  // |b|^2 - 2 a.b (+inf: padding)
                        v[i4 * 4 + j] = __float_as_uint(dd);
                        top2(j, dd);
                        if (DBG) { if (q < nq && i4 * 4 + j < rows) dbg[(size_t)q * nt + c * per + i4 * 4 + j] = __fadd_rn(na, dd); }
                    }
                }
                const float thr = merged_thr();
                // survivors are rare: branch-free bit mask per 32 columns, then walk the set bits
#pragma unroll
                for (int cb = 0; cb < NB; cb++) {
                    unsigned mask = 0u;
#pragma unroll
                    for (int i = 0; i < 32; i++) mask |= (__uint_as_float(v[cb * 32 + i]) <= thr ? 1u : 0u) << i;
                    while (mask) {
                        const int i = __ffs(mask) - 1;
                        mask &= mask - 1;
                        if (cn < kCand) s_cand[cn * kResThreads + tid] = (unsigned char)(cb * 32 + i);
                        cn++;
                    }
                }
            }
            store_rec(c, cn);
        }
    } else {
        for (int c = c_begin; c < c_end; c++) {
            const int rows = chunk_rows(c);
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // cp.async / st.shared -> visible to the tensor core
            __syncthreads();
            if (warp == 0 && rows > 0) {
                if (elect_one()) issue_mma(c);
                __syncwarp();
            }
            if (c + 1 < c_end) fetch(c + 1);  // the other tile buffer: its MMA (chunk c-1) completed before the last barrier
            int cn = 0;
            if (rows > 0) {
                mbar_wait(bar, phase);
                phase ^= 1;
                asm volatile("tcgen05.fence::after_thread_sync;");
                const float* nbp = s_nb + ((c - c_begin) % 3) * TRP;
                if (scans) {
#pragma unroll
                    for (int k = 0; k < 4; k++) { m1[k] = INFINITY; m2[k] = INFINITY; }
                    float thr = 0.f;
                    for (int pass = 0; pass < 2; pass++) {
                        for (int cb = 0; cb * 32 < rows; cb++) {
                            uint32_t v[32];
                            TMEM_LD32(v, taddr + cb * 32);
                            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                            const float4* nb4 = reinterpret_cast<const float4*>(nbp + cb * 32);
#pragma unroll
                            for (int j4 = 0; j4 < 8; j4++) {
                                const float4 nb = nb4[j4];
                                const float nbv[4] = {nb.x, nb.y, nb.z, nb.w};
#pragma unroll
                                for (int j = 0; j < 4; j++) {
                                    const float dd = __fadd_rn(__uint_as_float(v[j4 * 4 + j]), nbv[j]);
                                    if (pass == 0) {
                                        top2(j, dd);
                                        if (DBG) { if (q < nq && cb * 32 + j4 * 4 + j < rows) dbg[(size_t)q * nt + c * per + cb * 32 + j4 * 4 + j] = __fadd_rn(na, dd); }
                                    } else if (dd <= thr) {
                                        if (cn < kCand) s_cand[cn * kResThreads + tid] = (unsigned char)(cb * 32 + j4 * 4 + j);
                                        cn++;
                                    }
                                }
                            }
                        }
                        thr = merged_thr();
                    }
                }
            }
            store_rec(c, cn);
            asm volatile("tcgen05.fence::before_thread_sync;");
            __syncthreads();  // the accumulator is free again
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
    }
}

// ---------------------------------------------------------------- streaming filter (per > 256: large train sets)
// One CTA = 256 queries x ONE column chunk of the reference, streamed as 128-row train tiles, twice:
//   pass 0  an upper bound of the chunk's second smallest distance per query (branch-free: the 2nd smallest of the
//           minima of column octets is >= the true 2nd smallest, and equal to it unless both fall into one octet);
//   pass 1  the same tiles again (MMA recomputed: cheaper than keeping 2 x 256 x per distances anywhere), every
//           column with D <= thr(bound) is a survivor; an octet is only opened when its minimum passes.
// SINGLE = true is a one-pass variant (b200sift_match_set_path(3)): per step the running bound is updated first, then
// the same octets are tested against the threshold of the UPDATED running bound (never below the final one, so every
// final survivor is recorded when it is scanned) and passing columns go to a per-thread candidate list with their
// distance, re-filtered when more than half full and by the final threshold at the end -- the same survivors at half
// the MMA / TMA work. It is exact (same tests pass) but SLOWER: 5.69 vs 5.08 ms at 100k x 100k (profiles/r2_match_stream.md).
// With 32 independent query rows per warp some lane has a candidate in most octets while the bound is still loose, so
// the scan of a step costs ~2.3x, and the scan warps, not the tensor pipe, bound a step.
// Measured on B200 (tools/probes/tmem_probe.cu, elimination runs and ncu source views of earlier versions of this
// kernel): with both operands in shared memory a M128 N128 K16 MMA takes ~83 cycles instead of 64 (8 KB of operand
// reads per instruction saturate the 128 B/cycle shared-memory port), and streaming the train tiles with cp.async
// bounded every version at 9-10 ms for 100k x 100k whatever the number of warps issuing it (2176 sixteen-byte copies
// per tile, each holding its address registers until the memory system takes it). So:
//   * the train tiles come in by TMA (one elected thread, no registers, no issue slots): two 128-row x 64-half boxes
//     with the 128-byte swizzle (K-major SWIZZLE_128B operand layout) + one 4 KB bulk copy of the |b|^2 K chunk
//     (match_aug_kernel), all three completing on the tile's mbarrier; four tile buffers, filled two steps ahead;
//   * the QUERY tiles live in tensor memory (A operand from TMEM, written once per CTA with tcgen05.st: lane =
//     query row, 32-bit column j = halves 2j, 2j+1, pre-multiplied by -2): an MMA reads only its 4 KB train slice;
//   * K is 144, not 128: the ninth K step multiplies {2^15, 2^4, 2^-6} (query side) with the three-half split of
//     |b|^2, so the accumulator IS |b|^2 - 2 a.b and the scan needs no addition and no norm loads;
//   * 16 scan warps (lane quarter x query half x column half: 64 distances per thread and step, drained from TMEM
//     into registers in one go) + one warp whose lane 0 issues the TMA loads and the MMAs. The MMA of step s+1 runs
//     under the scan of step s.
// TMEM columns: [0, 256) accumulators (query half h at 128 h), [256, 400) query tiles (half h at 256 + 72 h).
// Tile buffer: [0, 16 KB) K 0..63 and [16, 32 KB) K 64..127, each [128 rows][128 B] swizzled; [32, 36 KB) the |b|^2
// chunk and a zero chunk, [2][128 rows][16 B] unswizzled.
constexpr int kStreamTN = 128;                         // train rows per tile = TMEM columns per accumulator
constexpr uint32_t kStreamPanel = 128 * 128;           // one swizzled K panel: 128 rows x 64 halves
constexpr uint32_t kStreamTile = 2 * kStreamPanel + 4096;
constexpr int kStreamBufs = 4;
constexpr int kStreamScan = 512;                       // 16 scan warps: (lane quarter, query half, column half)
constexpr int kStreamThreads = kStreamScan + 64;       // + one warp that issues the MMAs and one that issues the TMA loads
constexpr uint32_t kStreamAcol = 256;                  // first TMEM column of the query tiles
constexpr uint32_t kStreamAcols = 72;                  // 32-bit columns per query tile (144 halves)
constexpr int kRunCap = 16;                            // candidate list per scan thread (column + distance)
constexpr size_t kStreamSmem = kStreamBufs * (size_t)kStreamTile + 128 + (size_t)kRunCap * kStreamScan * (sizeof(int) + sizeof(float)) +
                               kStreamScan * sizeof(int) + 2 * kStreamScan * sizeof(float);
constexpr float kAugMaxNorm = 2.0e9f;                  // |b|^2 above this is not representable by the three-half split

#define TMEM_ST32(addr, v)                                                                                     \
    asm volatile(                                                                                              \
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "                                                        \
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "                             \
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"                     \
        ::"r"(addr), "r"((v)[0]), "r"((v)[1]), "r"((v)[2]), "r"((v)[3]), "r"((v)[4]), "r"((v)[5]), "r"((v)[6]), "r"((v)[7]),        \
          "r"((v)[8]), "r"((v)[9]), "r"((v)[10]), "r"((v)[11]), "r"((v)[12]), "r"((v)[13]), "r"((v)[14]), "r"((v)[15]),             \
          "r"((v)[16]), "r"((v)[17]), "r"((v)[18]), "r"((v)[19]), "r"((v)[20]), "r"((v)[21]), "r"((v)[22]), "r"((v)[23]),           \
          "r"((v)[24]), "r"((v)[25]), "r"((v)[26]), "r"((v)[27]), "r"((v)[28]), "r"((v)[29]), "r"((v)[30]), "r"((v)[31]) : "memory")

// K-major SWIZZLE_128B operand: rows of 128 B, 8-row groups 1024 B apart (SBO); the K step inside the 128-byte
// swizzle atom is taken by advancing the start address (32 B per 16 halves); tiles are 1024-byte aligned.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;                    // LBO: unused by swizzled K-major layouts
    d |= (uint64_t)(1024 >> 4) << 32;          // SBO
    d |= (uint64_t)1 << 46;                    // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
    return d;
}

template <bool SINGLE>
__global__ void __launch_bounds__(kStreamThreads, 1)
match_filter_stream_kernel(const __half* __restrict__ des, int nq, const __grid_constant__ CUtensorMap train_map, int nt, int per,
                           CandRec* __restrict__ rec, const float* __restrict__ nrm_q, const float* __restrict__ nrm_t,
                           const uint4* __restrict__ aug) {
    constexpr int TN = kStreamTN;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sB = smem;                                     // 4 tile buffers
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(sB + kStreamBufs * kStreamTile);  // [0..1] MMA done (query half), [2..5] tile landed, [6..9] tile buffer free
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 12);
    int* s_nbmax = reinterpret_cast<int*>(s_tmem + 1);
    int* s_cand = reinterpret_cast<int*>(s_bar + 16);       // [kRunCap][512] candidate columns (chunk-relative) of thread tid; the first kCand rows end up holding the survivors
    float* s_dist = reinterpret_cast<float*>(s_cand + kRunCap * kStreamScan);  // [kRunCap][512] their distances (|b|^2 - 2 a.b domain)
    int* s_cn = reinterpret_cast<int*>(s_dist + kRunCap * kStreamScan);        // [512] survivor counts
    float* s_pair = reinterpret_cast<float*>(s_cn + kStreamScan);  // [2][512]: (min, 2nd min) of every thread's columns

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool scanner = warp < kStreamScan / 32;
    const int half_id = (warp >> 2) & 1;                    // which 128-query tile
    const int col_half = (warp >> 3) & 1;                   // which 64 columns of every 128-column tile
    const int q0 = blockIdx.x * kResQ;
    const int row = half_id * kTileM + (warp & 3) * 32 + lane;  // query row inside the CTA (TMEM lane = row & 127)
    const int q = q0 + row;
    const int c = blockIdx.y;
    const int t_begin = c * per;
    const int crow = max(0, min(nt, t_begin + per) - t_begin);   // columns of this chunk
    if (crow == 0) {  // uniform for the CTA
        if (scanner && col_half == 0 && q < nq) rec[(size_t)q * 32 + c].n = 0;
        return;
    }
    const int T = (crow + TN - 1) / TN;
    const int tiles = (per + TN - 1) / TN;                  // tiles per chunk in the layout of `aug`
    constexpr uint32_t idesc = make_idesc(TN);
    const bool two_tiles = q0 + kTileM < nq;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        s_nbmax[0] = 0;
        for (int i = 0; i < 2 + 2 * kStreamBufs; i++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(s_bar + i)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = *s_tmem;
    if (scanner) {  // largest |b|^2 of the chunk (threshold slack); norms are >= 0: int order == float order
        float mx = 0.f;
        for (int t = tid; t < crow; t += kStreamScan) mx = fmaxf(mx, __ldg(nrm_t + t_begin + t));
        atomicMax(s_nbmax, __float_as_int(mx));
        // this thread's half of its query row (64 halves, x -2: exact) -> 32 TMEM columns of the row's lane; the
        // column-half-1 thread appends K chunk 16 = the multipliers of the |b|^2 split and the zero chunk 17
        uint32_t w[32];
        if (q < nq) {
            const uint4* g = reinterpret_cast<const uint4*>(des + (size_t)q * 128 + col_half * 64);
            const __half2 m2 = __floats2half2_rn(-2.f, -2.f);
#pragma unroll
            for (int i = 0; i < 8; i++) {
                uint4 x = __ldg(g + i);
                __half2* h = reinterpret_cast<__half2*>(&x);
#pragma unroll
                for (int j = 0; j < 4; j++) h[j] = __hmul2(h[j], m2);
                w[4 * i] = x.x; w[4 * i + 1] = x.y; w[4 * i + 2] = x.z; w[4 * i + 3] = x.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; i++) w[i] = 0u;
        }
        const uint32_t arow = tmem + ((uint32_t)((warp & 3) * 32) << 16) + kStreamAcol + (uint32_t)half_id * kStreamAcols;
        TMEM_ST32(arow + col_half * 32, w);
        if (col_half == 1) {
            const uint32_t m01 = (uint32_t)__half_as_ushort(__float2half_rn(kAugS1)) | ((uint32_t)__half_as_ushort(__float2half_rn(kAugS2)) << 16);
            const uint32_t m2_ = (uint32_t)__half_as_ushort(__float2half_rn(kAugS3));
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                         ::"r"(arow + 64), "r"(m01), "r"(m2_), "r"(0u), "r"(0u), "r"(0u), "r"(0u), "r"(0u), "r"(0u) : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    const float na = (scanner && q < nq) ? __ldg(nrm_q + q) : 0.f;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)half_id * TN + (uint32_t)col_half * (TN / 2);
    const uint32_t bar = smem_u32(s_bar);
    const float nbmax = __int_as_float(s_nbmax[0]);
    // |b|^2 beyond the three-half split (or not finite): every record of this CTA asks for the exact whole-chunk scan
    const bool bail = !(nbmax <= kAugMaxNorm);
    const int S = bail ? 0 : (SINGLE ? T : 2 * T);   // one pass over the chunk's tiles (two in the round-1 variant)

    const int nh = two_tiles ? 2 : 1;
    // step s works on tile s % T -> buffer s % 4; rows past the train set are zero-filled by the TMA unit, rows past
    // the chunk are real rows of the next chunk whose |b|^2 chunk says +inf
    auto load_tile = [&](int s) {
        const int ti = s >= T ? s - T : s;
        const uint32_t dst = smem_u32(sB) + (uint32_t)(s % kStreamBufs) * kStreamTile;
        const uint32_t full = bar + 8u * (2 + s % kStreamBufs);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"(kStreamTile) : "memory");
#pragma unroll
        for (int p = 0; p < 2; p++)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(dst + p * kStreamPanel), "l"(&train_map), "r"(p * 64), "r"(t_begin + ti * TN), "r"(full) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst + 2 * kStreamPanel), "l"(aug + ((size_t)c * tiles + ti) * 256), "r"(4096u), "r"(full) : "memory");
    };
    auto issue_mma = [&](int s, int hq) {  // query tile hq (TMEM) x the train tile of step s (shared memory) -> accumulator hq
        if (hq == 0) mbar_wait(bar + 8u * (2 + s % kStreamBufs), (uint32_t)(s / kStreamBufs) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t aB = smem_u32(sB) + (uint32_t)(s % kStreamBufs) * kStreamTile;
#pragma unroll
        for (int k = 0; k < 9; k++) {  // K = 144 = 9 x 16 halves = 9 x 8 TMEM columns of A
            const uint32_t ta = tmem + kStreamAcol + hq * kStreamAcols + k * 8;
            const uint64_t db = k < 8 ? umma_desc_sw128(aB + (k >> 2) * kStreamPanel + (k & 3) * 32)
                                      : umma_desc(aB + 2 * kStreamPanel, 128 * 16, 128);
            const uint32_t acc = k > 0 ? 1u : 0u;
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                ::"r"(tmem + hq * TN), "r"(ta), "l"(db), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar + 8u * hq) : "memory");
        if (hq == nh - 1)  // the step's last MMAs: once they are done the tile buffer can be refilled
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                         ::"r"(bar + 8u * (2 + kStreamBufs + s % kStreamBufs)) : "memory");
    };
    const float slack = __fmaf_rn(4e-5f, __fadd_rn(na, nbmax), 1e-3f);
    const bool scans = scanner && (half_id == 0 || two_tiles);
    int cn = 0;

    if (warp == kStreamScan / 32 + 1) {
        // ---- producer warp: the TMA loads have their own issuing thread. (Measured with clock64 stamps: issued from the
        // MMA thread they queue behind its UTCHMMAs and cost it ~390 of the ~1700 cycles of a step, during which no
        // further MMAs are queued.) Tile s goes into buffer s % 4 once the MMAs of tile s - 4 have completed.
        if (S > 0 && elect_one()) {
            for (int s = 0; s < S; s++) {
                if (s >= kStreamBufs) mbar_wait(bar + 8u * (2 + kStreamBufs + s % kStreamBufs), (uint32_t)(s / kStreamBufs - 1) & 1u);
                load_tile(s);
            }
        }
        __syncwarp();
    } else if (!scanner) {
        // ---- issuer warp: per step and query half h, wait until the scan warps of that half have drained their
        // accumulator (named barrier 1 + h: 256 arrivals + this warp), then queue the 9 MMAs of the next step behind
        // whatever the tensor pipe is still working on. Draining half 0 overlaps the MMAs of half 1 and vice versa.
        for (int s = 0; s < S; s++) {
            for (int h = 0; h < nh; h++) {
                if (s > 0) asm volatile("barrier.cta.sync %0, %1;" ::"r"(1 + h), "r"(kStreamScan / 2 + 32) : "memory");
                if (elect_one()) issue_mma(s, h);
                __syncwarp();
            }
        }
        if (S > 0)
            for (int h = 0; h < nh; h++) asm volatile("barrier.cta.sync %0, %1;" ::"r"(1 + h), "r"(kStreamScan / 2 + 32) : "memory");
    } else if (scans) {
      if constexpr (SINGLE) {
        // ---- scan warps (single pass, see the kernel's header comment)
        float m1[4] = {INFINITY, INFINITY, INFINITY, INFINITY}, m2[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
        const uint32_t done = bar + 8u * half_id;
        int rn = 0;          // candidates in this thread's list
        bool ovf = false;    // more than kRunCap columns within the current threshold: the exact path scores the whole chunk
        // {a1,a2} u {b1,b2}: min = min(a1,b1), 2nd = min(max(a1,b1), min(a2,b2))
        auto merged = [&](float& a1, float& a2) {
            const float x2 = fminf(fmaxf(m1[0], m1[1]), fminf(m2[0], m2[1])), x1 = fminf(m1[0], m1[1]);
            const float y2 = fminf(fmaxf(m1[2], m1[3]), fminf(m2[2], m2[3])), y1 = fminf(m1[2], m1[3]);
            a1 = fminf(x1, y1);
            a2 = fminf(fmaxf(x1, y1), fminf(x2, y2));
        };
        // Survivors: D <= D2nd * (1+d)/(1-d) with d = 1.5e-3 (half roundings) + fp32 accumulation slack, compared in the
        // |b|^2 - 2 a.b domain
        auto threshold = [&](float a2) {
            const float d2nd = fmaxf(__fadd_rn(a2, na), 0.f);
            return __fsub_rn(__fmaf_rn(d2nd, 1.004f, slack), na);
        };
        auto refilter = [&](float thr) {   // keep the candidates still within `thr`, order preserved
            int k = 0;
            for (int i = 0; i < rn; i++) {
                const float d = s_dist[i * kStreamScan + tid];
                if (d <= thr) {
                    s_dist[k * kStreamScan + tid] = d;
                    s_cand[k * kStreamScan + tid] = s_cand[i * kStreamScan + tid];
                    k++;
                }
            }
            rn = k;
        };
        for (int s = 0; s < S; s++) {
            uint32_t v[TN / 2];
            mbar_wait(done, (uint32_t)s & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;");
#pragma unroll
            for (int cb = 0; cb < TN / 64; cb++) TMEM_LD32(v + cb * 32, taddr + cb * 32);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;");
            asm volatile("barrier.cta.arrive %0, %1;" ::"r"(1 + half_id), "r"(kStreamScan / 2 + 32) : "memory");
            float om[TN / 16];
#pragma unroll
            for (int i8 = 0; i8 < TN / 16; i8++) {
                const float* f = reinterpret_cast<const float*>(v + i8 * 8);
                om[i8] = fminf(fminf(fminf(f[0], f[1]), fminf(f[2], f[3])), fminf(fminf(f[4], f[5]), fminf(f[6], f[7])));  // +inf on padding columns
                m2[i8 & 3] = fminf(m2[i8 & 3], fmaxf(m1[i8 & 3], om[i8]));
                m1[i8 & 3] = fminf(m1[i8 & 3], om[i8]);
            }
            float a1, a2;
            merged(a1, a2);
            const float thr = threshold(a2);   // from this thread's columns so far: >= the final threshold
            // 32 lanes = 32 independent query rows: in most octets SOME lane has a candidate, so this path is hot at warp
            // level. It is kept small: ONE list clean-up site per step (a list more than half full is re-filtered with
            // the current threshold, which has tightened since its entries were recorded), pushes are two predicated
            // stores, and a push that finds the list full only raises the overflow flag (-> exact whole-chunk scan).
            if (rn > kRunCap / 2) refilter(thr);
            if (!(thr < INFINITY)) ovf = true;   // fewer than two finite octets so far: cannot bound anything (never happens for nt > 8192)
            const int col0 = s * TN + col_half * (TN / 2);
#pragma unroll
            for (int i8 = 0; i8 < TN / 16; i8++) {
                if (om[i8] <= thr) {
                    const float* f = reinterpret_cast<const float*>(v + i8 * 8);
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        if (f[j] <= thr) {
                            const int slot = min(rn, kRunCap - 1) * kStreamScan + tid;
                            s_cand[slot] = col0 + i8 * 8 + j;
                            s_dist[slot] = f[j];
                            ovf |= rn >= kRunCap;
                            rn = min(rn + 1, kRunCap);
                        }
                    }
                }
            }
        }
        if (S > 0) {
            // the other column half of the row (thread tid ^ 256, same query half: named barrier 3 + half) -> final threshold
            float a1, a2v;
            merged(a1, a2v);
            s_pair[tid] = a1;
            s_pair[kStreamScan + tid] = a2v;
            asm volatile("barrier.cta.sync %0, %1;" ::"r"(3 + half_id), "r"(kStreamScan / 2) : "memory");
            const float b1 = s_pair[tid ^ 256], b2v = s_pair[kStreamScan + (tid ^ 256)];
            const float a2 = fminf(fmaxf(a1, b1), fminf(a2v, b2v));
            refilter(threshold(a2));
            cn = ovf ? 2 * kCand + 2 : rn;   // an overflowed list makes the record ask for the whole-chunk scan
        }
      } else {
        // ---- round-1 two-pass scan (kept for A/B measurements: b200sift_match_set_path(3))
        // ---- scan warps. Running (min, 2nd min) over the minima of column octets, four independent pairs: the 2nd
        // smallest octet minimum is >= the true 2nd smallest distance (equal unless both fall into one octet)
        float m1[4] = {INFINITY, INFINITY, INFINITY, INFINITY}, m2[4] = {INFINITY, INFINITY, INFINITY, INFINITY};
        float thr = 0.f;
        const uint32_t done = bar + 8u * half_id;
        for (int s = 0; s < S; s++) {
            uint32_t v[TN / 2];
            mbar_wait(done, (uint32_t)s & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;");
#pragma unroll
            for (int cb = 0; cb < TN / 64; cb++) TMEM_LD32(v + cb * 32, taddr + cb * 32);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;");
            asm volatile("barrier.cta.arrive %0, %1;" ::"r"(1 + half_id), "r"(kStreamScan / 2 + 32) : "memory");
            if (s < T) {
#pragma unroll
                for (int i8 = 0; i8 < TN / 16; i8++) {
                    const float* f = reinterpret_cast<const float*>(v + i8 * 8);
                    const float om = fminf(fminf(fminf(f[0], f[1]), fminf(f[2], f[3])), fminf(fminf(f[4], f[5]), fminf(f[6], f[7])));  // +inf on padding columns
                    m2[i8 & 3] = fminf(m2[i8 & 3], fmaxf(m1[i8 & 3], om));
                    m1[i8 & 3] = fminf(m1[i8 & 3], om);
                }
                if (s == T - 1) {
                    // {a1,a2} u {b1,b2}: min = min(a1,b1), 2nd = min(max(a1,b1), min(a2,b2)); then the same with the other
                    // column half of the row (thread tid ^ 256, same query half: named barrier 3 + half).
                    // Survivors: D <= D2nd * (1+d)/(1-d) with d = 1.5e-3 (half roundings) + fp32 accumulation slack,
                    // compared in the |b|^2 - 2 a.b domain
                    const float x2 = fminf(fmaxf(m1[0], m1[1]), fminf(m2[0], m2[1])), x1 = fminf(m1[0], m1[1]);
                    const float y2 = fminf(fmaxf(m1[2], m1[3]), fminf(m2[2], m2[3])), y1 = fminf(m1[2], m1[3]);
                    const float a1 = fminf(x1, y1), a2v = fminf(fmaxf(x1, y1), fminf(x2, y2));
                    s_pair[tid] = a1;
                    s_pair[kStreamScan + tid] = a2v;
                    asm volatile("barrier.cta.sync %0, %1;" ::"r"(3 + half_id), "r"(kStreamScan / 2) : "memory");
                    const float b1 = s_pair[tid ^ 256], b2v = s_pair[kStreamScan + (tid ^ 256)];
                    const float a2 = fminf(fmaxf(a1, b1), fminf(a2v, b2v));
                    const float d2nd = fmaxf(__fadd_rn(a2, na), 0.f);
                    thr = __fsub_rn(__fmaf_rn(d2nd, 1.004f, slack), na);
                }
            } else {
                const int col0 = (s - T) * TN + col_half * (TN / 2);
#pragma unroll
                for (int i8 = 0; i8 < TN / 16; i8++) {
                    const float* f = reinterpret_cast<const float*>(v + i8 * 8);
                    const float om = fminf(fminf(fminf(f[0], f[1]), fminf(f[2], f[3])), fminf(fminf(f[4], f[5]), fminf(f[6], f[7])));
                    if (om <= thr) {  // survivors are rare
#pragma unroll
                        for (int j = 0; j < 8; j++) {
                            if (f[j] <= thr) {
                                if (cn < kCand) s_cand[cn * kStreamScan + tid] = col0 + i8 * 8 + j;
                                cn++;
                            }
                        }
                    }
                }
            }
        }
      }
    }
    if (scanner) s_cn[tid] = cn;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (scanner && col_half == 0 && q < nq) {
        // merge the two column halves of the row in ascending column order (both lists already are)
        const int pt = tid ^ 256;
        const int na_ = min(cn, kCand), nb_ = min(s_cn[pt], kCand);
        const int total = bail ? kCand + 1 : cn + s_cn[pt];
        int cc[kCand];
        int ia = 0, ib = 0;
#pragma unroll
        for (int i = 0; i < kCand; i++) {
            const int ca = ia < na_ ? s_cand[ia * kStreamScan + tid] : 0x7fffffff;
            const int cb = ib < nb_ ? s_cand[ib * kStreamScan + pt] : 0x7fffffff;
            const int m = min(ca, cb);
            cc[i] = (m == 0x7fffffff || bail) ? -1 : t_begin + m;
            if (ca <= cb) ia++; else ib++;
        }
        int4* dst = reinterpret_cast<int4*>(rec + (size_t)q * 32 + c);
        dst[0] = make_int4(min(total, kCand + 1), cc[0], cc[1], cc[2]);
        dst[1] = make_int4(cc[3], cc[4], cc[5], cc[6]);
    }
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
    }
}

// The reference's per-pair score (Match.cu:93-111) is warp-cooperative: lane i owns half2 i and i+32 of both rows and a
// shfl_down tree (offsets 16, 8, 4, 2, 1) sums the 32 partial sums. The re-rank kernel below plays it with 8 lanes per
// pair (four pairs per warp step): the same operand pairs are added in the same tree, fp32 addition is commutative, so
// every score is bit-identical to the reference's lane-0 value.
constexpr int kRerankWarps = 4;
constexpr int kListCap = 32 * kCand;

__global__ void __launch_bounds__(kRerankWarps * 32)
match_rerank_kernel(const __half* __restrict__ des, int nq, const __half* __restrict__ src, int nt, int per,
                    const CandRec* __restrict__ rec, int* __restrict__ out_idx, const int* __restrict__ cnt_q,
                    const int* __restrict__ cnt_t) {
    const int max_q = nq;
    if (cnt_q) {
        nt = *cnt_t;
        nq = *cnt_q;
        per = (int)ceilf((float)nt / 32.f);
    }
    __shared__ int s_col[kRerankWarps][kListCap];
    __shared__ float s_val[kRerankWarps][kListCap];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * kRerankWarps + warp;
    if (q >= nq) {  // device-side sizes: the grid covers max_n rows, rows past the query count report "no match"
        if (q < max_q && lane == 0) out_idx[q] = -1;
        return;
    }
    const __half2* arow = reinterpret_cast<const __half2*>(des + (size_t)q * 128);
    // lane c <- survivors of chunk c
    const int4* rp = reinterpret_cast<const int4*>(rec + (size_t)q * 32 + lane);
    const int4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
    const int cnt = r0.x;
    const bool overflow = cnt > kCand;
    const int mine = overflow ? 0 : cnt;
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += u;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    {
        const int base = incl - mine;
        const int c[kCand] = {r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
        for (int i = 0; i < kCand; i++)
            if (i < mine) s_col[warp][base + i] = c[i];
    }
    __syncwarp();
    float mn = 1e6f, mn2 = 1e6f;   // Match.cu:125-129 sentinels (quirk Q16)
    int id = -1, id2 = -1;
    // Survivors, ascending column order inside every chunk, FOUR per step: lane 8g+l works on survivor p+g and
    // plays the reference's lanes l, l+8, l+16, l+24 (half2 l+8k, k = 0..7, of both rows). The reference's
    // reduction tree (Match.cu:93-111: offsets 16, 8, 4, 2, 1) pairs lane v with v+16, then v+8: those two
    // levels are local adds here, the last three are xor shuffles inside the 8-lane group, so every sum has
    // the same operands in the same tree and the score is bit-identical.
    {
        const int g = lane >> 3, l = lane & 7;
        __half2 a8[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a8[k] = arow[l + 8 * k];
        const __half2 inv = __floats2half2_rn(0.25f, 0.25f);
        auto fetch = [&](int p, __half2* b8) {
            const int col = s_col[warp][min(p + g, total - 1)];
            const __half2* brow = reinterpret_cast<const __half2*>(src + (size_t)col * 128);
#pragma unroll
            for (int k = 0; k < 8; k++) b8[k] = __ldg(brow + l + 8 * k);
        };
        __half2 b8[8], n8[8];
        if (total > 0) fetch(0, b8);
        for (int p = 0; p < total; p += 4) {
            if (p + 4 < total) fetch(p + 4, n8);
            float sv[4];  // partial sums of the reference's lanes l, l+8, l+16, l+24
#pragma unroll
            for (int v = 0; v < 4; v++) {
                __half2 d0 = __hmul2(__hsub2(a8[v], b8[v]), inv);          // half2 index l+8v      (reference lane's first)
                __half2 d1 = __hmul2(__hsub2(a8[v + 4], b8[v + 4]), inv);  // half2 index l+8v+32   (its second)
                d0 = __hmul2(d0, d0);
                d1 = __hmul2(d1, d1);
                const float2 f0 = __half22float2(d0), f1 = __half22float2(d1);
                sv[v] = __fadd_rn(__fadd_rn(f0.x, f0.y), __fadd_rn(f1.x, f1.y));
            }
            float val = __fadd_rn(__fadd_rn(sv[0], sv[2]), __fadd_rn(sv[1], sv[3]));  // offsets 16 then 8
#pragma unroll
            for (int off = 4; off > 0; off >>= 1) val = __fadd_rn(val, __shfl_xor_sync(0xffffffffu, val, off));
            if (l == 0 && p + g < total) s_val[warp][p + g] = val;
#pragma unroll
            for (int k = 0; k < 8; k++) b8[k] = n8[k];
        }
        __syncwarp();
        // lane c replays its own chunk's survivors in list (= ascending column) order, strict '<' as the reference
        const int base = incl - mine;
        for (int i = 0; i < mine; i++) {
            const float v = s_val[warp][base + i];
            const int col = s_col[warp][base + i];
            if (v < mn) { mn2 = mn; id2 = id; mn = v; id = col; }
            else if (v < mn2) { mn2 = v; id2 = col; }
        }
    }
    // overflowed chunks: every column of the chunk, exactly as the reference scans them. Sixteen columns per step:
    // lane group g (8 lanes) scores column t + 4s + g with the same 8-lane arithmetic as above (bit-identical to the
    // reference's tree), all 32 loads of a step are issued before the first score, and lane `ch` applies the 16 results
    // in ascending column order. (One column per step, as this loop first was, made a single overflowed chunk of a
    // 100k-row train set a 0.5 ms serial chain: the floor of the whole kernel, whatever the number of query rows.)
    unsigned ovf = __ballot_sync(0xffffffffu, overflow);
    if (ovf) {
        const int g = lane >> 3, l = lane & 7;
        __half2 a8[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a8[k] = arow[l + 8 * k];
        const __half2 inv = __floats2half2_rn(0.25f, 0.25f);
        while (ovf) {
            const int ch = __ffs(ovf) - 1;
            ovf &= ovf - 1;
            const int t_end = min(nt, (ch + 1) * per);
            for (int t = ch * per; t < t_end; t += 16) {
                __half2 bb[4][8];
#pragma unroll
                for (int st = 0; st < 4; st++) {
                    const int col = min(t + 4 * st + g, t_end - 1);
                    const __half2* brow = reinterpret_cast<const __half2*>(src + (size_t)col * 128);
#pragma unroll
                    for (int k = 0; k < 8; k++) bb[st][k] = __ldg(brow + l + 8 * k);
                }
#pragma unroll
                for (int st = 0; st < 4; st++) {
                    float sv[4];
#pragma unroll
                    for (int v = 0; v < 4; v++) {
                        __half2 d0 = __hmul2(__hsub2(a8[v], bb[st][v]), inv);
                        __half2 d1 = __hmul2(__hsub2(a8[v + 4], bb[st][v + 4]), inv);
                        d0 = __hmul2(d0, d0);
                        d1 = __hmul2(d1, d1);
                        const float2 f0 = __half22float2(d0), f1 = __half22float2(d1);
                        sv[v] = __fadd_rn(__fadd_rn(f0.x, f0.y), __fadd_rn(f1.x, f1.y));
                    }
                    float val = __fadd_rn(__fadd_rn(sv[0], sv[2]), __fadd_rn(sv[1], sv[3]));
#pragma unroll
                    for (int off = 4; off > 0; off >>= 1) val = __fadd_rn(val, __shfl_xor_sync(0xffffffffu, val, off));
#pragma unroll
                    for (int gg = 0; gg < 4; gg++) {
                        const float vv = __shfl_sync(0xffffffffu, val, 8 * gg);
                        const int col = t + 4 * st + gg;
                        if (lane == ch && col < t_end) {
                            if (vv < mn) { mn2 = mn; id2 = id; mn = vv; id = col; }
                            else if (vv < mn2) { mn2 = vv; id2 = col; }
                        }
                    }
                }
            }
        }
    }
    // 32-lane merge, same shuffle sequence as Match.cu:152-169 (second update uses the partner's min2
    // after the partner's own first update: quirk Q15 reproduced by construction)
    float min_dist = mn, min_dist_2 = mn2;
    int min_idx = id, min_idx_2 = id2;
#pragma unroll
    for (int offset = 16; offset > 0; offset /= 2) {
        float nearby_min = __shfl_down_sync(0xffffffffu, min_dist, offset);
        float nearby_min_idx = (float)__shfl_down_sync(0xffffffffu, min_idx, offset);
        if (nearby_min < min_dist) {
            min_dist_2 = min_dist;
            min_idx_2 = min_idx;
            min_dist = nearby_min;
            min_idx = (int)nearby_min_idx;
        }
        nearby_min = __shfl_down_sync(0xffffffffu, min_dist_2, offset);
        nearby_min_idx = (float)__shfl_down_sync(0xffffffffu, min_idx_2, offset);
        if (nearby_min < min_dist_2) {
            min_dist_2 = nearby_min;
            min_idx_2 = (int)nearby_min_idx;
        }
    }
    if (lane == 0) out_idx[q] = ((double)min_dist < 0.8 * (double)min_dist_2) ? min_idx : -1;
}

}  // namespace

namespace {
constexpr int kMaxDevices = 64;
// Device the calling thread launches on (index into the per-device attribute tables below). One thread per GPU may
// run concurrently: every table entry is only ever touched by the threads whose current device it belongs to, and
// setting an attribute twice is harmless.
int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < kMaxDevices) ? dev : 0;
}
int g_match_path = 0;  // 0 = tensor-core filter + exact re-rank, 1 = all-CUDA-core exact kernel (match.cu), 2 = as 0 with the unpipelined generic filter for large train sets
constexpr size_t filter_smem(int tn) { return kTileBytes + (size_t)tn * 256 + tn * sizeof(float) + 64 + kCand * kFilterThreads * sizeof(int); }

template <int TN, bool DBG>
void launch_filter(dim3 grid, const __half* q, int nq, const __half* t, int nt, int per, CandRec* rec, float* dbg, cudaStream_t s,
                   const int* cnt_q = nullptr, const int* cnt_t = nullptr) {
    // the attribute belongs to the (device, context) pair, not to the process: track it per device
    static size_t smem_set[kMaxDevices] = {0};
    const int dev = current_device();
    if (smem_set[dev] < filter_smem(TN)) {
        cudaFuncSetAttribute(match_filter_kernel<TN, DBG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)filter_smem(TN));
        smem_set[dev] = filter_smem(TN);
    }
    match_filter_kernel<TN, DBG><<<grid, kFilterThreads, filter_smem(TN), s>>>(q, nq, t, nt, per, rec, dbg, cnt_q, cnt_t);
}
}  // namespace

namespace {
// Tensor map of the train set for the streaming filter's TMA loads: [nt rows][128 halves], box = 128 rows x 64
// halves with the 128-byte swizzle; rows past nt read as zero. cuTensorMapEncodeTiled comes from the driver through
// the runtime's entry-point query (no link-time dependency on libcuda).
bool encode_train_map(CUtensorMap* map, const __half* t, int nt) {
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
    const cuuint64_t dims[2] = {128, (cuuint64_t)nt};
    const cuuint64_t strides[1] = {256};
    const cuuint32_t box[2] = {64, (cuuint32_t)kStreamTN};
    const cuuint32_t estr[2] = {1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half*>(t), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
}  // namespace

void match_set_path(int path) { g_match_path = path; }
int match_get_path() { return g_match_path; }

size_t match_scratch_bytes(int nq, int nt) {
    // survivor records (>= the 16-byte records of the CUDA-core path) + the squared norms of both sets
    // + the |b|^2 K chunks of the streaming filter (4 KB per 128-row tile of every chunk)
    const int per = (nt + 31) / 32;
    const size_t aug = per > 256 ? (size_t)32 * ((per + kStreamTN - 1) / kStreamTN) * 4096 : 0;
    return (size_t)nq * 32 * sizeof(CandRec) + ((size_t)nq + (size_t)nt + 8) * sizeof(float) + aug;
}

int match_num_launches(int nq, int nt) { return (nq > 0 && nt >= 0) ? (nt > 8192 ? 4 : 3) : 0; }  // norms, (|b|^2 chunks,) filter, re-rank

namespace {
// Resident-chunk filter (every chunk of the train set fits one MMA tile): norms, then the filter.
template <bool DBG, int NB>
void launch_resident_nb(dim3 grid, size_t smem, const __half* q, int nq, const __half* t, int nt, int per, int TR, int cpc, CandRec* rec,
                        float* dbg, const float* nrm_q, const float* nrm_t, cudaStream_t s, const int* cnt_q, const int* cnt_t) {
    static size_t smem_set[kMaxDevices] = {0};
    const int dev = current_device();
    if (smem > smem_set[dev]) {
        cudaFuncSetAttribute(match_filter_resident_kernel<DBG, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        smem_set[dev] = smem;
    }
    match_filter_resident_kernel<DBG, NB><<<grid, kResThreads, smem, s>>>(q, nq, t, nt, per, TR, cpc, rec, dbg, nrm_q, nrm_t, cnt_q, cnt_t);
}

template <bool DBG>
void launch_resident(const __half* q, int nq, const __half* t, int nt, int per, CandRec* rec, float* dbg, cudaStream_t s,
                     const int* cnt_q, const int* cnt_t) {
    float* nrm_q = reinterpret_cast<float*>(rec + (size_t)nq * 32);
    float* nrm_t = nrm_q + nq;
    match_norms_kernel<<<(nq + nt + 7) / 8, 256, 0, s>>>(q, nq, t, nt, nrm_q, nrm_t, cnt_q, cnt_t);
    const int TR = std::max(32, (per + 31) & ~31);
    // at least 116 KB so that one CTA (which owns all 512 TMEM columns) is resident per SM
    const size_t smem = std::max<size_t>(2 * kTileBytes + (size_t)2 * TR * 256 + (size_t)3 * TR * 4 + 64 + kCand * kResThreads, 116 * 1024);
    const int tiles = (nq + kResQ - 1) / kResQ;
    int cpc = 8;  // chunks per CTA: as many as still leave about one CTA per SM
    while (cpc > 1 && tiles * (32 / cpc) < 140) cpc >>= 1;
    dim3 grid(tiles, 32 / cpc);
#define LAUNCH_NB(NB) launch_resident_nb<DBG, NB>(grid, smem, q, nq, t, nt, per, TR, cpc, rec, dbg, nrm_q, nrm_t, s, cnt_q, cnt_t)
    switch (TR / 32) {
        case 1: LAUNCH_NB(1); break;
        case 2: LAUNCH_NB(2); break;
        case 3: LAUNCH_NB(3); break;
        case 4: LAUNCH_NB(4); break;
        case 5: LAUNCH_NB(5); break;
        default: LAUNCH_NB(0); break;
    }
#undef LAUNCH_NB
}
}  // namespace

int match_launch(const __half* q, int nq, const __half* t, int nt, int* out, void* scratch, size_t scratch_bytes,
                 cudaStream_t s, float* dbg, cudaEvent_t* marks) {
    // marks (profiling only): [0] after the norm (+ |b|^2 chunk) kernels, [1] after the filter, [2] after the re-rank
    auto mark = [&](int i) { if (marks) cudaEventRecord(marks[i], s); };
    if (nq <= 0) return 0;
    if (scratch_bytes < match_scratch_bytes(nq, nt)) return -1;
    if (g_match_path == 1 && !dbg) return match_launch_simt(q, nq, t, nt, out, scratch, scratch_bytes, s);
    CandRec* rec = static_cast<CandRec*>(scratch);
    const int per = (int)ceilf((float)nt / 32.f);  // Match.cu:131 (float arithmetic)
    if (per <= 256) {
        if (dbg) launch_resident<true>(q, nq, t, nt, per, rec, dbg, s, nullptr, nullptr);
        else launch_resident<false>(q, nq, t, nt, per, rec, dbg, s, nullptr, nullptr);
    } else if (dbg || g_match_path == 2) {
        dim3 grid((nq + kTileM - 1) / kTileM, 32);
        if (dbg) launch_filter<256, true>(grid, q, nq, t, nt, per, rec, dbg, s);
        else launch_filter<256, false>(grid, q, nq, t, nt, per, rec, dbg, s);
    } else {
        static bool attr_set[kMaxDevices] = {false};
        const int dev = current_device();
        if (!attr_set[dev]) {
            cudaFuncSetAttribute(match_filter_stream_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStreamSmem);
            cudaFuncSetAttribute(match_filter_stream_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStreamSmem);
            attr_set[dev] = true;
        }
        // scratch: records | |b|^2 K chunks, tiled per chunk (4 KB per 128-row tile) | norms
        const int tiles = (per + kStreamTN - 1) / kStreamTN;
        uint4* aug = reinterpret_cast<uint4*>(rec + (size_t)nq * 32);
        float* nrm_q = reinterpret_cast<float*>(aug + (size_t)32 * tiles * 256);
        float* nrm_t = nrm_q + nq;
        CUtensorMap train_map;
        if (!encode_train_map(&train_map, t, nt)) return -2;
        match_norms_kernel<<<(nq + nt + 7) / 8, 256, 0, s>>>(q, nq, t, nt, nrm_q, nrm_t, nullptr, nullptr);
        match_aug_kernel<<<(32 * tiles * 128 + 255) / 256, 256, 0, s>>>(nrm_t, nt, per, tiles, aug);
        mark(0);
        dim3 grid((nq + kResQ - 1) / kResQ, 32);
        // single-pass variant (path 3): measured 5.69 ms against 5.08 ms for the two passes at 100k x 100k on one B200 (r2): a
        // step is bounded by the scan warps, not by the tensor pipe, and one pass makes every step's scan more expensive
        if (g_match_path == 3) match_filter_stream_kernel<true><<<grid, kStreamThreads, kStreamSmem, s>>>(q, nq, train_map, nt, per, rec, nrm_q, nrm_t, aug);
        else match_filter_stream_kernel<false><<<grid, kStreamThreads, kStreamSmem, s>>>(q, nq, train_map, nt, per, rec, nrm_q, nrm_t, aug);
    }
    if (per <= 256 || dbg || g_match_path == 2) mark(0);  // those paths compute their norms inside / right before the filter
    mark(1);
    match_rerank_kernel<<<(nq + kRerankWarps - 1) / kRerankWarps, kRerankWarps * 32, 0, s>>>(q, nq, t, nt, per, rec, out, nullptr, nullptr);
    mark(2);
    return 0;
}

// Same matcher with the set sizes read on the device: *cnt_q = query rows, *cnt_t = train rows, both
// <= max_n. Grids are sized for max_n; surplus CTAs exit at once.
int match_launch_counts(const __half* q, const __half* t, int max_n, const int* cnt_q, const int* cnt_t, int* out, void* scratch,
                        size_t scratch_bytes, cudaStream_t s) {
    if (max_n <= 0) return 0;
    if (scratch_bytes < match_scratch_bytes(max_n, max_n)) return -1;
    CandRec* rec = static_cast<CandRec*>(scratch);
    const int per_max = (max_n + 31) / 32;
    if (per_max <= 256) {
        launch_resident<false>(q, max_n, t, max_n, per_max, rec, nullptr, s, cnt_q, cnt_t);
    } else {
        dim3 grid((max_n + kTileM - 1) / kTileM, 32);
        launch_filter<256, false>(grid, q, max_n, t, max_n, per_max, rec, nullptr, s, cnt_q, cnt_t);
    }
    match_rerank_kernel<<<(max_n + kRerankWarps - 1) / kRerankWarps, kRerankWarps * 32, 0, s>>>(q, max_n, t, max_n, per_max, rec, out, cnt_q, cnt_t);
    return 0;
}

}  // namespace b200sift
