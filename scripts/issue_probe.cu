// issue_probe.cu — exploratory micro-benchmark (not part of the product): how many SMSP cycles does one warp-iteration of
// the escape-time loop cost on B200 for different instruction mixes around the 6 FP64 instructions?
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -o issue_probe issue_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

struct O { double x, y, a, b, cr, ci; };

__device__ __forceinline__ void core6(O &o) {
    const double s = __dsub_rn(o.a, o.b);
    const double p = __dmul_rn(o.x, o.y);
    const double nre = __dadd_rn(s, o.cr);
    const double nim = __fma_rn(p, 2.0, o.ci);
    o.x = nre; o.y = nim;
    o.a = __dmul_rn(nre, nre);
    o.b = __dmul_rn(nim, nim);
}
__device__ __forceinline__ void core7(O &o) {
    const double s = __dsub_rn(o.a, o.b);
    const double t = __dadd_rn(o.x, o.x);
    const double nim = __dadd_rn(__dmul_rn(t, o.y), o.ci);
    const double nre = __dadd_rn(s, o.cr);
    o.x = nre; o.y = nim;
    o.a = __dmul_rn(nre, nre);
    o.b = __dmul_rn(nim, nim);
}
__device__ __forceinline__ bool filt(const O &o, uint32_t T) {
    return max((uint32_t)__double2hiint(o.a), (uint32_t)__double2hiint(o.b)) >= T;
}
__device__ __forceinline__ O init(int k) {
    O o; o.cr = -0.1 + 1e-4 * (threadIdx.x + k); o.ci = 0.1 + 1e-5 * (blockIdx.x % 97);
    o.x = o.cr; o.y = o.ci; o.a = o.x * o.x; o.b = o.y * o.y; return o;
}

// MODE 0: core6 only. 1: core6 + filter+branch each iter. 2: core6 + filter every 2nd iter (max of both).
// 3: core6 + one VIMNMX accumulate (no branch). 4: core7 + filter. 5: core6 x2 orbits + filter each. 6: core6 + filter every 4th
// 7: core6 + ISETP.OR predicate accumulate, branch every 8.
template <int MODE>
__global__ void __launch_bounds__(256, 4) probe(double *sink, int trips, uint32_t T)
{
    O o = init(0), o2 = init(7);
    uint32_t acc = 0; bool esc = false;
    for (int t = 0; t < trips && !esc; ++t) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (MODE == 4) core7(o); else core6(o);
            if (MODE == 5) core6(o2);
            if (MODE == 1 || MODE == 4 || MODE == 5) { if (filt(o, T)) { esc = true; break; } }
            if (MODE == 5) { if (filt(o2, T)) { esc = true; break; } }
            if (MODE == 2) { acc = max(acc, max((uint32_t)__double2hiint(o.a), (uint32_t)__double2hiint(o.b))); if (u & 1) { if (acc >= T) { esc = true; break; } } }
            if (MODE == 6) { acc = max(acc, max((uint32_t)__double2hiint(o.a), (uint32_t)__double2hiint(o.b))); if ((u & 3) == 3) { if (acc >= T) { esc = true; break; } } }
            if (MODE == 3) { acc = max(acc, (uint32_t)__double2hiint(o.a)); }
            if (MODE == 7) { acc = max(acc, max((uint32_t)__double2hiint(o.a), (uint32_t)__double2hiint(o.b))); }
        }
        if (MODE == 7 || MODE == 3) { if (acc >= T) esc = true; }
    }
    if (esc || o.x == 123.0 || o2.x == 77.0) sink[0] = o.x + o2.y + acc;
}

template <int MODE>
void run(const char *name, double *sink, int blocks_per_sm)
{
    int dev; cudaGetDevice(&dev); cudaDeviceProp pr; cudaGetDeviceProperties(&pr, dev);
    const int trips = 4000;
    const int blocks = pr.multiProcessorCount * blocks_per_sm;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<blocks, 256>>>(sink, 100, 0x41e00000u);
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0);
        probe<MODE><<<blocks, 256>>>(sink, trips, 0x41e00000u);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double orbits = (MODE == 5) ? 2.0 : 1.0;
    const double warp_iters_per_smsp = (double)blocks_per_sm * 8 /*warps*/ / 4.0 * trips * 8 * orbits;
    const double cycles = best * 1e-3 * 1.965e9;
    const double iters_per_s = (double)blocks * 256 * trips * 8 * orbits / (best * 1e-3);
    printf("%-44s blocks/SM=%d  %8.3f ms  %6.2f cyc/warp-iter  %7.1f G iter/s\n", name, blocks_per_sm, best, cycles / warp_iters_per_smsp, iters_per_s / 1e9);
}

int main()
{
    double *sink; cudaMalloc(&sink, 64);
    for (int bps : {2, 4, 5}) {
        run<0>("core6 only", sink, bps);
        run<1>("core6 + filter(VIMNMX,ISETP,BRA) every iter", sink, bps);
        run<2>("core6 + VIMNMX x2, ISETP+BRA every 2nd", sink, bps);
        run<6>("core6 + VIMNMX x2, ISETP+BRA every 4th", sink, bps);
        run<7>("core6 + VIMNMX x2, ISETP+BRA every 8th", sink, bps);
        run<3>("core6 + 1 VIMNMX, check every 8th", sink, bps);
        run<4>("core7 + filter every iter", sink, bps);
        run<5>("2 orbits/thread: core6 + filter every iter", sink, bps);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
