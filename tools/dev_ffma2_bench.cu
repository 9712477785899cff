// Development microbenchmark: scalar FFMA (register x uniform-register tap) against packed FFMA2 (fma.rn.f32x2) on sm_100a:
// same number of IEEE fp32 FMAs, half the instructions. nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/ffma2 tools/dev_ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
struct Taps { float v[16]; float2 v2[16]; };
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b), rc = *reinterpret_cast<unsigned long long*>(&c), rd;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
__global__ void __launch_bounds__(256) k_scalar(float* out, const __grid_constant__ Taps t, int iters) {
    float a[16];
#pragma unroll
    for (int j = 0; j < 16; j++) a[j] = threadIdx.x * 0.001f + j;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++)
#pragma unroll
            for (int j = 0; j < 16; j++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[j]) : "f"(t.v[k]), "f"(a[(j + 1) & 15]));
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) s += a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) k_packed(float* out, const __grid_constant__ Taps t, int iters) {
    float2 a[8];
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = make_float2(threadIdx.x * 0.001f + j, threadIdx.x * 0.002f + j);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++)
#pragma unroll
            for (int j = 0; j < 8; j++) a[j] = fma2(a[j], t.v2[k], a[(j + 1) & 7]);
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s += a[j].x + a[j].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    Taps t;
    for (int i = 0; i < 16; i++) { t.v[i] = 0.5f + 0.01f * i; t.v2[i] = make_float2(t.v[i], t.v[i]); }
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 2000;
    for (int rep = 0; rep < 3; rep++) {
        for (int which = 0; which < 2; which++) {
            cudaEventRecord(e0);
            if (which == 0) k_scalar<<<148 * 8, 256>>>(out, t, iters); else k_packed<<<148 * 8, 256>>>(out, t, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double fmas = 148.0 * 8 * 256 * iters * 256;  // scalar FMAs (per thread per iteration: 256)
            printf("%s: %.3f ms, %.2f TFMA/s (%.1f TFLOP/s), %s\n", which ? "FFMA2 packed" : "FFMA scalar", ms, fmas / ms / 1e9, 2 * fmas / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
        }
    }
    return 0;
}
