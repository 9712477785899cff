// Micro-probe (development tool, not product): TMEM drain rate (tcgen05.ld 32x32b.x32) for 4/8/16 warps, the
// SS-mode tcgen05.mma rate at M=128 N=128 K=16 with the no-swizzle K-major layout, and both at once.
// Build + run on the GPU box:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmem_probe tools/probes/tmem_probe.cu && /tmp/tmem_probe
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
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

// mode bit 0: warps drain TMEM (each warp 128 columns = 4 x32 loads per iteration); bit 1: thread 0 of the LAST warp issues MMAs
// bit 5 (32): the warp before the MMA warp streams 32 KB bulk copies global -> shared memory for as long as the MMA warp runs;
// bit 6 (64): the B operand is read through K-major SWIZZLE_128B descriptors (as the matcher does) instead of the no-swizzle layout
__global__ void __launch_bounds__(1024) probe(int mode, int iters, int mma_per_iter, long long* out, unsigned* sink, const uint4* gsrc) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ uint64_t s_bar, s_bar2, s_bar3;
    __shared__ volatile int s_stop;
    __shared__ int s_copies;
    const int tid = threadIdx.x, warp = tid >> 5, nwarps = blockDim.x >> 5;
    for (int i = tid; i < 3 * 36864 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar2)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar3)));
        s_stop = 0; s_copies = 0;
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = s_tmem;
    const bool mma_warp = (mode & 2) && warp == nwarps - 1;
    const bool drains = (mode & 1) && !((mode & 2) && warp == nwarps - 1);
    const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) & 3) * 128;
    unsigned acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    if (mma_warp) {
        uint32_t el;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(el));
        if (el) {
        const int n = (mode & 8) ? 64 : 128;
        const uint32_t idesc = (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        uint32_t phase = 0;
        for (int it = 0; it < iters; it++) {
            for (int g = 0; g < mma_per_iter / 9; g++) {
#pragma unroll
                for (int k = 0; k < 9; k++) {
                    const uint64_t da = umma_desc(smem_u32(smem) + k * 2 * 2048, 2048, 128);
                    uint64_t db = umma_desc(smem_u32(smem) + 36864 + k * 2 * 2048, 2048, 128);
                    if ((mode & 64) && k < 8) {
                        const uint32_t a = smem_u32(smem) + 36864 + (k >> 2) * 16384 + (k & 3) * 32;
                        db = (uint64_t)((a & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
                    }
                    const uint32_t accf = k > 0 ? 1u : 0u;
                    if (mode & 4)
                        asm volatile(
                            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                            ::"r"(tmem + (g & 1) * 128), "r"(tmem + 256 + k * 8), "l"(db), "r"(idesc), "r"(accf) : "memory");
                    else
                        asm volatile(
                            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                            ::"r"(tmem + (g & 1) * 128), "l"(da), "l"(db), "r"(idesc), "r"(accf) : "memory");
                }
                // mode bit 4: a commit after every group of 9 (onto a second barrier nobody waits for), as the matcher does
                if ((mode & 16) && g + 1 < mma_per_iter / 9)
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_bar2)) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_bar)) : "memory");
            uint32_t done = 0;
            while (!done) {
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"(smem_u32(&s_bar)), "r"(phase) : "memory");
            }
            phase ^= 1;
        }
        s_stop = 1;
        }
        __syncwarp();
    } else if ((mode & 32) && warp == nwarps - 2) {
        uint32_t el;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(el));
        if (el) {
            uint32_t ph = 0; int n = 0;
            while (!s_stop) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&s_bar3)), "r"(32768u) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(smem) + 2 * 36864), "l"(gsrc + (size_t)(n & 255) * 2048), "r"(32768u), "r"(smem_u32(&s_bar3)) : "memory");
                uint32_t done = 0;
                while (!done)
                    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                                 : "=r"(done) : "r"(smem_u32(&s_bar3)), "r"(ph) : "memory");
                ph ^= 1; n++;
            }
            s_copies = n;
        }
        __syncwarp();
    } else if (drains) {
        for (int it = 0; it < iters; it++) {
            uint32_t v[128];
#pragma unroll
            for (int cb = 0; cb < 4; cb++) TMEM_LD32(v + cb * 32, taddr + cb * 32);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < 128; i += 8) acc ^= v[i];
        }
    }
    const long long t1 = clock64();
    if ((tid & 31) == 0 && blockIdx.x == 0) out[warp] = t1 - t0;
    __syncthreads();
    if (tid == 0 && blockIdx.x == 0) out[63] = s_copies;
    if (acc == 0x12345678u) sink[0] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

int main() {
    long long* d_out; unsigned* d_sink;
    cudaMalloc(&d_out, 64 * sizeof(long long)); cudaMalloc(&d_sink, 4);
    uint4* d_src; cudaMalloc(&d_src, 256 * 32768); cudaMemset(d_src, 0, 256 * 32768);
    const size_t smem = 3 * 36864;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 2000;
    struct Cfg { int mode, warps, mma; const char* name; } cfgs[] = {
        {1, 4, 0, "drain only, 4 warps (128 rows x 128 cols per iter)"},
        {1, 8, 0, "drain only, 8 warps (2 x 128 x 128 per iter)"},
        {1, 16, 0, "drain only, 16 warps (4 x 128 x 128 per iter)"},
        {2, 1, 18, "mma only, 18 x (M128 N128 K16) per commit"},
        {2, 1, 36, "mma only, 36 per commit"},
        {3, 9, 18, "8 drain warps + mma 18 per commit"},
        {3, 5, 18, "4 drain warps + mma 18 per commit"},
        {6, 1, 18, "mma A-from-TMEM N=128, 18 per commit"},
        {6, 1, 36, "mma A-from-TMEM N=128, 36 per commit"},
        {14, 1, 36, "mma A-from-TMEM N=64, 36 per commit"},
        {14, 1, 72, "mma A-from-TMEM N=64, 72 per commit"},
        {7, 9, 18, "8 drain warps + mma A-from-TMEM N=128, 18 per commit"},
        {6, 1, 72, "mma A-from-TMEM N=128, 72 per wait, one commit"},
        {22, 1, 72, "mma A-from-TMEM N=128, 72 per wait, commit every 9"},
        {23, 9, 72, "same + 8 drain warps"},
        {6 + 64, 1, 72, "mma A-from-TMEM N=128, 72 per wait, B through SWIZZLE_128B"},
        {6 + 32, 2, 72, "mma A-from-TMEM N=128 (no-swizzle B) + bulk-copy stream"},
        {6 + 64 + 32, 2, 72, "mma A-from-TMEM N=128, SWIZZLE_128B B + bulk-copy stream"},
        {7 + 64 + 32, 10, 72, "same + 8 drain warps"},
    };
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int grid : {1}) {
        for (auto& c : cfgs) {
            const int it = grid == 1 ? iters : iters * 20;  // long enough for the power management to settle
            cudaMemset(d_out, 0, 64 * sizeof(long long));
            cudaEventRecord(e0);
            probe<<<grid, c.warps * 32, smem>>>(c.mode, it, c.mma, d_out, d_sink, d_src);
            cudaEventRecord(e1);
            cudaError_t e = cudaDeviceSynchronize();
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            long long h[64];
            cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
            printf("grid %3d  %-52s err=%s  cycles/iter: warp0 %.1f  last warp %.1f   %.2f ms -> %.0f MHz\n", grid, c.name, cudaGetErrorString(e),
                   (double)h[0] / it, (double)h[c.warps - 1] / it, ms, (double)h[c.warps - 1] / (ms * 1e3));
            if (c.mode & 32) printf("          bulk copies: %lld x 32 KB = %.1f B/cycle\n", h[63], (double)h[63] * 32768 / (double)h[c.warps - 1]);
        }
    }
    return 0;
}
