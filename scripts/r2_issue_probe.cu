// Does an FP64 instruction leave its second pipe cycle free for another instruction of the same SM sub-partition?
// Loop body: 6 dependent-chain FP64 instructions (the iteration) + N integer instructions (independent of the chain), 16 warps per
// sub-partition.  If cycles/iteration = max(12, 6 + N) the issue port is free during the FP64 pipe's second cycle; if it is
// 12 + N an FP64 instruction costs two issue slots.   nvcc -O3 -arch=sm_100a -o r2_issue_probe r2_issue_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int N>
__global__ void __launch_bounds__(512) probe(double *out, unsigned *iout, int iters, long long *cycles)
{
    double x = 0.1 + threadIdx.x * 1e-3, y = 0.2, a = x * x, b = y * y, cr = -0.7, ci = 0.1;
    unsigned u0 = threadIdx.x, u1 = blockIdx.x, u2 = 7, u3 = 9;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            double s, t;
            asm volatile("sub.rn.f64 %0, %1, %2;" : "=d"(s) : "d"(a), "d"(b));
            asm volatile("mul.rn.f64 %0, %1, %2;" : "=d"(t) : "d"(x), "d"(y));
            asm volatile("add.rn.f64 %0, %1, %2;" : "=d"(x) : "d"(s), "d"(cr));
            asm volatile("fma.rn.f64 %0, %1, 0d4000000000000000, %2;" : "=d"(y) : "d"(t), "d"(ci));
            asm volatile("mul.rn.f64 %0, %1, %1;" : "=d"(a) : "d"(x));
            asm volatile("mul.rn.f64 %0, %1, %1;" : "=d"(b) : "d"(y));
            if (N >= 1) asm volatile("xor.b32 %0, %0, %1;" : "+r"(u0) : "r"(u1));
            if (N >= 2) asm volatile("add.u32 %0, %0, %1;" : "+r"(u1) : "r"(u2));
            if (N >= 3) asm volatile("xor.b32 %0, %0, %1;" : "+r"(u2) : "r"(u3));
            if (N >= 4) asm volatile("add.u32 %0, %0, %1;" : "+r"(u3) : "r"(u0));
            if (N >= 5) asm volatile("xor.b32 %0, %0, %1;" : "+r"(u0) : "r"(u2));
            if (N >= 6) asm volatile("add.u32 %0, %0, %1;" : "+r"(u1) : "r"(u3));
            if (N >= 7) asm volatile("xor.b32 %0, %0, %1;" : "+r"(u2) : "r"(u0));
            if (N >= 8) asm volatile("add.u32 %0, %0, %1;" : "+r"(u3) : "r"(u1));
            if (N >= 9) asm volatile("xor.b32 %0, %0, %1;" : "+r"(u0) : "r"(u3));
            if (N >= 10) asm volatile("add.u32 %0, %0, %1;" : "+r"(u1) : "r"(u0));
            if (N >= 11) asm volatile("xor.b32 %0, %0, %1;" : "+r"(u2) : "r"(u1));
            if (N >= 12) asm volatile("add.u32 %0, %0, %1;" : "+r"(u3) : "r"(u2));
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x + y + a + b;
    iout[blockIdx.x * blockDim.x + threadIdx.x] = u0 + u1 + u2 + u3;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int N>
void run(double *out, unsigned *iout, long long *cyc)
{
    const int iters = 4096;
    for (int warps_per_smsp : {2, 4, 8, 16}) {
        probe<N><<<148, 128 * warps_per_smsp / 1 > 1024 ? 1024 : 128 * warps_per_smsp, 0>>>(out, iout, iters, cyc);
        cudaDeviceSynchronize();
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        const int threads = 128 * warps_per_smsp > 1024 ? 1024 : 128 * warps_per_smsp;
        const int blocks = 148 * (128 * warps_per_smsp / threads);
        cudaEventRecord(e0);
        probe<N><<<blocks, threads>>>(out, iout, iters, cyc);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        const double per_iter_warp = (double)c / (iters * 4.0) / warps_per_smsp;
        printf("N=%2d warps/SMSP=%2d: %.2f cycles per warp-iteration (issue-port cycles per iteration; 12 = FP64 pipe bound)  kernel %.3f ms\n", N,
               warps_per_smsp, per_iter_warp, ms);
    }
}

int main()
{
    double *out; unsigned *iout; long long *cyc;
    cudaMalloc(&out, 148 * 2 * 1024 * 8); cudaMalloc(&iout, 148 * 2 * 1024 * 4); cudaMalloc(&cyc, 8);
    run<0>(out, iout, cyc); run<2>(out, iout, cyc); run<4>(out, iout, cyc); run<6>(out, iout, cyc); run<8>(out, iout, cyc); run<12>(out, iout, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
