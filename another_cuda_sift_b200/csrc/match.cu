// Brute-force ratio-test matcher on CUDA cores, bit-exact with sift_func/Match.cu:35-177. This is the
// on-device cross-check of the tensor-core path in match_tc.cu (b200sift_match_set_path(1)).
//
// The reference computes, for query q and train t,
//     score = fp32 tree-sum over 32 lanes of ((e[2i] + e[2i+1]) + (e[2i+64] + e[2i+65])),
//     e[j]  = half( d*d ),  d = half( half(a[j]-b[j]) * 0.25 )                    (Match.cu:93-106)
// writes the full nq x nt score matrix to HBM, re-reads it so that lane i can take the top-2 of
// the contiguous column chunk [i*ceil(nt/32), ...), and merges the 32 lane results with a
// shuffle tree whose second-best update is order dependent (quirk Q15).
//
// Here the score matrix never exists. Kernel 1 gives every thread one (query, chunk) pair: the
// query row lives in registers, train rows are broadcast from shared memory, the 32 "lane"
// partial sums and the shuffle tree are evaluated in registers in the reference's exact order,
// and the chunk's top-2 is kept on the fly. Kernel 2 replays the reference's 32-lane merge with
// real warp shuffles (lane i <- chunk i), so Q15 is reproduced by construction, and applies the
// 0.8 ratio test in double (Match.cu:172).
#include "common.cuh"

namespace b200sift {

namespace {

constexpr int kMatchThreads = 128;
constexpr int kTrainTile = 16;

struct __align__(16) ChunkRec {
    float mn;
    int idx;
    float mn2;
    int idx2;
};

__global__ void __launch_bounds__(kMatchThreads)
match_chunks_kernel(const __half* __restrict__ des, int nq, const __half* __restrict__ src, int nt, int per,
                    ChunkRec* __restrict__ rec) {
    __shared__ __align__(16) __half2 s_train[kTrainTile][64];
    const int q = blockIdx.x * kMatchThreads + threadIdx.x;
    const int chunk = blockIdx.y;
    const int t_begin = chunk * per;
    const int t_end = min(nt, t_begin + per);
    const bool active = q < nq;

    __half2 a[64];
    {
        const uint4* p = reinterpret_cast<const uint4*>(des + (size_t)(active ? q : 0) * 128);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            uint4 u = __ldg(p + i);
            a[4 * i + 0] = *reinterpret_cast<__half2*>(&u.x);
            a[4 * i + 1] = *reinterpret_cast<__half2*>(&u.y);
            a[4 * i + 2] = *reinterpret_cast<__half2*>(&u.z);
            a[4 * i + 3] = *reinterpret_cast<__half2*>(&u.w);
        }
    }
    const __half2 inv = __floats2half2_rn(0.25f, 0.25f);
    float mn = 1e6f, mn2 = 1e6f;   // sentinels of Match.cu:125-129 (quirk Q16)
    int id = -1, id2 = -1;

    for (int t0 = t_begin; t0 < t_end; t0 += kTrainTile) {
        const int rows = min(kTrainTile, t_end - t0);
        __syncthreads();
        // stage `rows` train rows (256 B each) with 16-byte loads
        for (int i = threadIdx.x; i < rows * 16; i += kMatchThreads) {
            const int r = i >> 4, c = i & 15;
            reinterpret_cast<uint4*>(&s_train[r][0])[c] =
                __ldg(reinterpret_cast<const uint4*>(src + (size_t)(t0 + r) * 128) + c);
        }
        __syncthreads();
        if (!active) continue;
        for (int r = 0; r < rows; r++) {
            float s[32];
#pragma unroll
            for (int i = 0; i < 32; i++) {
                const __half2 b0 = s_train[r][i], b1 = s_train[r][i + 32];
                __half2 d0 = __hmul2(__hsub2(a[i], b0), inv);
                __half2 d1 = __hmul2(__hsub2(a[i + 32], b1), inv);
                d0 = __hmul2(d0, d0);
                d1 = __hmul2(d1, d1);
                const float2 f0 = __half22float2(d0), f1 = __half22float2(d1);
                s[i] = __fadd_rn(__fadd_rn(f0.x, f0.y), __fadd_rn(f1.x, f1.y));
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1)
#pragma unroll
                for (int i = 0; i < off; i++) s[i] = __fadd_rn(s[i], s[i + off]);
            const float val = s[0];
            const int t = t0 + r;
            if (val < mn) { mn2 = mn; id2 = id; mn = val; id = t; }
            else if (val < mn2) { mn2 = val; id2 = t; }
        }
    }
    if (active) {
        ChunkRec rc;
        rc.mn = mn; rc.idx = id; rc.mn2 = mn2; rc.idx2 = id2;
        rec[(size_t)q * 32 + chunk] = rc;
    }
}

// Warp per query; lane i holds the top-2 of chunk i. Same shuffle sequence as Match.cu:152-169:
// the second pair of shuffles reads the partner's min2 AFTER the partner's own first update.
__global__ void __launch_bounds__(256)
match_merge_kernel(const ChunkRec* __restrict__ rec, int nq, int* __restrict__ out_idx) {
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;
    const ChunkRec rc = rec[(size_t)q * 32 + lane];
    float min_dist = rc.mn, min_dist_2 = rc.mn2;
    int min_idx = rc.idx, min_idx_2 = rc.idx2;
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

int match_launch_simt(const __half* q, int nq, const __half* t, int nt, int* out, void* scratch, size_t scratch_bytes, cudaStream_t s) {
    if (nq <= 0) return 0;
    if (scratch_bytes < (size_t)nq * 32 * sizeof(ChunkRec)) return -1;
    ChunkRec* rec = static_cast<ChunkRec*>(scratch);
    // ceil(float(nt)/32) as in Match.cu:131 (float arithmetic)
    const int per = (int)ceilf((float)nt / 32.f);
    dim3 grid((nq + kMatchThreads - 1) / kMatchThreads, 32);
    match_chunks_kernel<<<grid, kMatchThreads, 0, s>>>(q, nq, t, nt, per, rec);
    match_merge_kernel<<<(nq + 7) / 8, 256, 0, s>>>(rec, nq, out);
    return 0;
}

}  // namespace b200sift
