// DFMA / F2F throughput probe (experiment, not product code)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) probe_dfma(double* out, int iters) {
    double x[8];
    for (int u = 0; u < 8; ++u) x[u] = threadIdx.x * 1e-3 + u;
    const double m = 0.999, c = 1e-3 * (blockIdx.x + 1);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int u = 0; u < 8; ++u) x[u] = fma(x[u], m, c);
    }
    double s = 0;
    for (int u = 0; u < 8; ++u) s += x[u];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    double* out;
    cudaMalloc(&out, sizeof(double) * 148 * 8 * 256);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    const int iters = 512, blocks = 148 * 8, threads = 256;
    probe_dfma<<<blocks, threads>>>(out, iters);
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    probe_dfma<<<blocks, threads>>>(out, iters);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    printf("DFMA: %.3f ms  %.2f TFLOP/s fp64\n", ms, 2.0 * 64 * iters * (double)blocks * threads / ms * 1e-9);
    return 0;
}
