// latency_probe.cu — exploratory micro-benchmark (not part of the product): cycles per iteration of the 6-instruction loop
// body as a function of the warps resident per SMSP.  One warp alone is bound by the dependent chain (DADD -> DADD -> DMUL);
// enough warps are bound by the FP64 pipe (6 instructions x 2 issue cycles).  The crossover says how many warps per SMSP the
// iterate kernel needs, and how fast a single late pixel runs once the rest of the frame has drained (the tail of a frame).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -o latency_probe latency_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

__global__ void probe(double *sink, long long *cyc, int iters)
{
    double cr = -0.1 + 1e-4 * threadIdx.x, ci = 0.1 + 1e-5 * (blockIdx.x % 97);
    double x = cr, y = ci, a = x * x, b = y * y;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 16
    for (int k = 0; k < iters; ++k) {
        const double s = __dsub_rn(a, b);
        const double q = __dmul_rn(x, y);
        x = __dadd_rn(s, cr);
        y = __fma_rn(q, 2.0, ci);
        a = __dmul_rn(x, x);
        b = __dmul_rn(y, y);
        // keep the orbit bounded without a branch the compiler can predict away: values only matter for timing
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    if (x == 123.456) sink[0] = x + y + a + b;
}

int main()
{
    int dev = 0;
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, dev);
    double *sink;
    long long *cyc;
    cudaMalloc(&sink, 64);
    cudaMalloc(&cyc, sizeof(long long) * 4096);
    const int iters = 20000;
    for (int threads : {32, 128, 256, 384, 512, 640, 768, 1024}) {
        probe<<<pr.multiProcessorCount, threads>>>(sink, cyc, 100);
        cudaDeviceSynchronize();
        probe<<<pr.multiProcessorCount, threads>>>(sink, cyc, iters);
        cudaDeviceSynchronize();
        long long h[4096];
        cudaMemcpy(h, cyc, sizeof(long long) * pr.multiProcessorCount, cudaMemcpyDeviceToHost);
        double sum = 0;
        for (int i = 0; i < pr.multiProcessorCount; ++i) sum += (double)h[i];
        const double per_iter = sum / pr.multiProcessorCount / iters;
        const double warps_per_smsp = threads / 32 / 4.0;
        printf("threads/SM %4d (%.2f warps/SMSP): %7.2f cycles per warp-iteration, pipe busy %.3f\n", threads, warps_per_smsp, per_iter,
               warps_per_smsp < 1 ? 12.0 / per_iter : warps_per_smsp * 12.0 / per_iter);
    }
    printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
