// frb_tile.cuh — K1t: the short-orbit kernel.  One thread per pixel, warps on 8x4-pixel tiles, the escape filter after every
// iteration, colour computed in the same kernel from the registers.  Included by frb_kernels.cuh.
//
// The lane-refill kernels (K1c and its predecessors) pay per PIXEL: a wasted half block, the idle rest of a trip, a share of a
// service call, a 20-byte record through HBM and up to B+1 lock-step resolve steps — about 30-50 SMSP cycles, which is
// nothing against C2's 1435 iterations per pixel and everything against C1's 41 (K1c: 0.30 of the FP64 peak there).  With
// a small iteration cap the lock-step alternative cannot lose much: a tile costs at most cap x 16 cycles, and measured tile
// utilisation is high exactly where orbits are short (C1, 8x4 tiles: 0.93 — SURVEY Appendix C).  So frames whose cap is at
// most kTileMaxIterations (frb_api.cu) take this kernel: per iteration the same 6 (or, where FORM 6 is not provably exact,
// 7) FP64 instructions as everywhere else plus VIMNMX / ISETP / BRA, no queue, no records, no second kernel.
// Same arithmetic, same exact test (calc/src/lib.rs:249-254 on the carried squares), so counts and final positions are the
// ones every other kernel produces; the raw-result path (frb_render_counts) keeps using those.
#pragma once

namespace frb {

constexpr int kTileW = 32, kTileH = 8;  // pixels per CTA: 8 warps, warp w covers columns (w % 4) * 8 .. + 8, rows (w / 4) * 4 .. + 4

template <int FORM>
__global__ void __launch_bounds__(kTileW * kTileH)
frb_iterate_tile_kernel(const __grid_constant__ RenderParams p)
{
    __shared__ __align__(16) uint8_t stage[kTileH][kTileW * 4];
    __shared__ double2 lg[128];
    __shared__ unsigned long long s_useful[kTileH];
    __shared__ uint32_t s_longest[kTileH];
    load_log2_table(lg);  // needed by shade_pixel only: the barrier that publishes it comes after the loop
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint32_t tx = (warp & 3u) * 8u + (lane & 7u), ty = (warp >> 2) * 4u + (lane >> 3);
    const uint32_t x = blockIdx.x * kTileW + tx;
    const uint32_t lr = blockIdx.y * kTileH + ty;  // row of this launch; p.resolve_row0 = its first row within the shard
    const bool valid = x < p.width && lr < p.rows;

    uint32_t count = 0;
    Orbit o = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    if (valid) {
        const double re0 = __ldg(p.tab_re + x);
        const double im0 = __ldg(p.tab_im + local_row_to_y(p, p.resolve_row0 + lr));
        const bool julia = (p.algo == kAlgoJulia);
        // get_recursive_pixel (calc/src/lib.rs:208-210): Mandelbrot c = start, Julia c = julia_set
        o = {re0, im0, __dmul_rn(re0, re0), __dmul_rn(im0, im0), julia ? p.julia_re : re0, julia ? p.julia_im : im0};
        const double S = p.limit_sq;
        const uint32_t filter_hi = p.filter_hi;
        const uint32_t iterations = p.iterations;
        count = iterations;
        // recursive (calc/src/lib.rs:245-257): (next, i) on escape, (previous, iterations) otherwise.  Four iterations per
        // loop trip; per iteration the 6 (7) FP64 instructions, VIMNMX + ISETP + BRA for the filter, and nothing else: the
        // exact test and the count live behind the (rarely taken) filter branch.
        uint32_t i = 0;
        // The exact test sits in a PTX loop that never repeats (i < iterations <= 2^32 - 1): a back edge is the one thing
        // ptxas will not if-convert, so DADD + DSETP run only behind the filter branch instead of on every iteration.
#define FRB_TILE_STEP(K)                                                                              \
    stream_core<FORM>(o);                                                                             \
    if (stream_filter(o, filter_hi)) {                                                                \
        uint32_t esc;                                                                                 \
        asm volatile("{\n\t.reg .pred q, r;\n\t.reg .f64 s;\n"                                         \
                     "FRB_TILE_%=:\n\t"                                                               \
                     "add.rn.f64 s, %1, %2;\n\t"                                                      \
                     "setp.gt.f64 q, s, %3;\n\t"                                                      \
                     "selp.u32 %0, 1, 0, q;\n\t"                                                      \
                     "setp.eq.u32 r, %4, 0xffffffff;\n\t"                                             \
                     "@r bra FRB_TILE_%=;\n\t}"                                                       \
                     : "=r"(esc)                                                                      \
                     : "d"(o.a), "d"(o.b), "d"(S), "r"(i));                                           \
        if (esc) {                                                                                    \
            count = i + (K);                                                                          \
            goto escaped;                                                                             \
        }                                                                                             \
    }
        for (const uint32_t n4 = iterations & ~3u; i < n4; i += 4) {
            FRB_TILE_STEP(0)
            FRB_TILE_STEP(1)
            FRB_TILE_STEP(2)
            FRB_TILE_STEP(3)
        }
        for (; i < iterations; ++i) {
            FRB_TILE_STEP(0)
        }
#undef FRB_TILE_STEP
    escaped:;
    }
    __syncthreads();
    const uint32_t c = valid ? shade_pixel(p, count, make_double2(o.x, o.y), lg) : 0u;

    // statistics: loop-body executions (i + 1 | iterations) and the lane slots the lock-step warps spent (32 x the longest orbit),
    // summed over the CTA first: two same-address atomics per WARP (65 000 warps on C1) serialise in the L2 for longer than
    // the frame takes to iterate
    {
        const uint32_t useful = (valid && p.iterations != 0u) ? min(count, p.iterations - 1u) + 1u : 0u;
        const uint32_t longest = __reduce_max_sync(0xffffffffu, useful);
        const uint32_t lo = __reduce_add_sync(0xffffffffu, useful & 0xffffu);  // two halves: 32 x 2^32 does not fit a word
        const uint32_t hi = __reduce_add_sync(0xffffffffu, useful >> 16);
        if (lane == 0) {
            s_useful[warp] = (unsigned long long)lo + ((unsigned long long)hi << 16);
            s_longest[warp] = longest;
        }
    }
    if (p.format != 4u) {
        stage[ty][tx * 3 + 0] = (uint8_t)c;
        stage[ty][tx * 3 + 1] = (uint8_t)(c >> 8);
        stage[ty][tx * 3 + 2] = (uint8_t)(c >> 16);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long useful = 0, slots = 0;
#pragma unroll
        for (int w = 0; w < kTileH; ++w) {
            useful += s_useful[w];
            slots += 32ull * s_longest[w];
        }
        if (useful != 0ull) {
            atomicAdd(p.stats + 0, useful);
            atomicAdd(p.stats + 1, slots);
        }
    }

    if (p.format == 4u) {
        if (valid) reinterpret_cast<uint32_t *>(p.out)[(size_t)lr * p.width + x] = c | 0xff000000u;
        return;
    }
    // RGB24: the CTA's 8 row segments of up to 96 bytes.  Full segments of a frame whose width is a multiple of 4 start on a
    // word boundary (row * width * 3 + 96 * blockIdx.x): 24 word stores per row, consecutive threads on consecutive words.
    const uint32_t x0 = blockIdx.x * kTileW;
    const uint32_t seg = min((uint32_t)kTileW, p.width - x0) * 3u;  // bytes per row segment
    if (seg == 96u && (p.width & 3u) == 0u && (reinterpret_cast<uintptr_t>(p.out) & 3u) == 0u) {
        if (threadIdx.x < (uint32_t)kTileH * 24u) {
            const uint32_t r = threadIdx.x / 24u, w = threadIdx.x - r * 24u;
            const uint32_t row = blockIdx.y * kTileH + r;
            if (row < p.rows)
                reinterpret_cast<uint32_t *>(p.out + ((size_t)row * p.width + x0) * 3u)[w] = reinterpret_cast<const uint32_t *>(stage[r])[w];
        }
        return;
    }
    for (uint32_t k = threadIdx.x; k < (uint32_t)kTileH * 96u; k += blockDim.x) {
        const uint32_t r = k / 96u, b = k - r * 96u;
        const uint32_t row = blockIdx.y * kTileH + r;
        if (b < seg && row < p.rows) p.out[((size_t)row * p.width + x0) * 3u + b] = stage[r][b];
    }
}

}  // namespace frb
