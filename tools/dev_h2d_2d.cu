// Development microbenchmark: host->device copy rate of one 1920x1200 float frame as cudaMemcpyAsync (1-D) against
// cudaMemcpy2DAsync with equal pitches (what the engine issued), pinned host memory, one stream.
#include <cstdio>
#include <cuda_runtime.h>
int main() {
    const int w = 1920, h = 1200, nb = 8, iters = 40;
    const size_t bytes = (size_t)w * h * 4;
    float *hbuf[nb], *dbuf[nb];
    for (int i = 0; i < nb; i++) { cudaHostAlloc(&hbuf[i], bytes, cudaHostAllocDefault); cudaMalloc(&dbuf[i], bytes); }
    cudaStream_t s; cudaStreamCreate(&s);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 4; mode++) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0, s);
            for (int it = 0; it < iters; it++)
                for (int i = 0; i < nb; i++) {
                    if (mode == 0) cudaMemcpyAsync(dbuf[i], hbuf[i], bytes, cudaMemcpyHostToDevice, s);
                    else if (mode == 1) cudaMemcpy2DAsync(dbuf[i], (size_t)w * 4, hbuf[i], (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyHostToDevice, s);
                    else if (mode == 2) cudaMemcpyAsync(hbuf[i], dbuf[i], bytes / 6, cudaMemcpyDeviceToHost, s);
                    else cudaMemcpy2DAsync(dbuf[i], (size_t)2048 * 4, hbuf[i], (size_t)w * 4, (size_t)(w - 128) * 4, h, cudaMemcpyHostToDevice, s);
                }
            cudaEventRecord(e1, s); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double b = mode == 2 ? bytes / 6 : (mode == 3 ? (double)(w - 128) * 4 * h : bytes);
            if (rep) printf("mode %d (%s): %.2f GB/s, %.1f us per copy\n", mode, mode == 0 ? "1-D H2D" : mode == 1 ? "2-D H2D equal pitches" : mode == 2 ? "1-D D2H 1.5 MB" : "2-D H2D padded rows",
                            b * iters * nb / ms / 1e6, ms * 1e3 / (iters * nb));
        }
    }
    return 0;
}
