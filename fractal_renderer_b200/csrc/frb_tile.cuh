// frb_tile.cuh — K1t: the short-orbit kernel.  One thread per pixel, independent warps on 8x4-pixel (or 32x1) tiles, one escape
// filter per pair of iterations, colour computed in the same kernel from the registers.  Included by frb_kernels.cuh.
//
// The lane-refill kernels (K1c and its predecessors) pay per PIXEL: a wasted half block, the idle rest of a trip, a share of a
// service call, a 20-byte record through HBM and up to B+1 lock-step resolve steps — about 30-50 SMSP cycles, which is
// nothing against C2's 1435 iterations per pixel and everything against C1's 41 (K1c: 0.30 of the FP64 peak there).  With
// a small iteration cap the lock-step alternative cannot lose much: a tile costs at most cap x 16 cycles, and measured tile
// utilisation is high exactly where orbits are short (C1, 8x4 tiles: 0.93 — SURVEY Appendix C).  So frames whose cap is at
// most kTileMaxIterations (frb_api.cu) take this kernel: per iteration the same 6 (or, where FORM 6 is not provably exact,
// 7) FP64 instructions as everywhere else plus the filter of a PAIR of iterations (VIMNMX, VIMNMX3, ISETP, BRA), no queue, no
// records, no second kernel.
// Same arithmetic, same exact test (calc/src/lib.rs:249-254 on the carried squares), so counts and final positions are the
// ones every other kernel produces; the raw-result path (frb_render_counts) keeps using those.
#pragma once

namespace frb {

constexpr uint32_t kTileStatsWord = 16u;   // first word of p.stats that holds the tile kernel's spread counters
constexpr uint32_t kTileStatsPairs = 32u;  // (useful iterations, lane slots) pairs

// TW x (32 / TW) pixels per warp, WPC warps side by side per CTA.  Warps share nothing — no shared memory, no barrier — so the
// CTA is only the granularity at which the SM gets warp slots back: with 8-warp CTAs and a barrier before the stores (the
// first version) a tile on the rim of the set kept seven finished warps resident behind its one long orbit, and the FP64
// pipe starved with the SM nominally full (ncu: 44 resident warps, pipe 61 % busy, 2.2 barrier stalls per issue).
template <int FORM, int TW, int WPC>
__global__ void __launch_bounds__(32 * WPC)
frb_iterate_tile_kernel(const __grid_constant__ RenderParams p)
{
    constexpr uint32_t TH = 32u / TW;
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t tx = lane % TW, ty = lane / TW;
    const uint32_t x0 = (blockIdx.x * WPC + (threadIdx.x >> 5)) * TW;  // first column of the warp's tile
    const uint32_t x = x0 + tx;
    const uint32_t lr = blockIdx.y * TH + ty;  // row of this launch; p.resolve_row0 = its first row within the shard
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
        const uint32_t n4 = iterations & ~3u;
        count = iterations;
        // recursive (calc/src/lib.rs:245-257): (next, i) on escape, (previous, iterations) otherwise.
        // The steady state is a loop of four iterations with ONE filter test per pair: max(hi(re^2), hi(im^2)) over both
        // positions of the pair on the integer pipe (VIMNMX, VIMNMX3), ISETP, BRA — 8.75 issue slots an iteration for 6 on
        // the FP64 pipe, which is what bounds the loop.  A pair computed past an escape is harmless: squares grow to inf or
        // NaN, whose high words also fire the filter, and only the positions up to the first exact escape are used.  A lane
        // whose filter fires LEAVES the loop and only outside it runs the exact tests, the count, and — when the filter fired
        // early (S/2 <= max(re^2, im^2) but re^2 + im^2 <= S) — the exact steps up to the next multiple of four before it goes
        // back in.  Leaving (instead of an `if` inside the body) keeps the convergence barrier (BSSY/BSYNC: two issue slots
        // an iteration in the first version of this kernel) out of the loop; the price is that a lane with an early hit
        // waits outside until the others leave as well, which costs nothing where it matters (limit^2 >= 8: the step after
        // an early hit escapes; and escaped lanes idle to the end of the tile anyway).
        uint32_t i = 0;
        for (;;) {
            Orbit pre;  // written by the loop on the exits that use it
            uint32_t k;
            // Written in PTX: from C++ the compiler recomputed squares on both sides of the filter branch and carried the
            // exit index through the loop (26 FP64 and 12 other instructions per four iterations instead of 24 and 11).
            // Exits: k = 4 with i == n4 (no hit), or k = 0 / 2 = first iteration of the pair that fired, `pre` = the position
            // after that iteration and `o` = the position after the second one.
#define FRB_STEP6(X, Y, A, B, NX, NY, NA, NB)                \
    "sub.rn.f64 s, " A ", " B ";\n\t"                        \
    "mul.rn.f64 t, " X ", " Y ";\n\t"                        \
    "add.rn.f64 " NX ", s, %10;\n\t"                         \
    "fma.rn.f64 " NY ", t, 0d4000000000000000, %11;\n\t"     \
    "mul.rn.f64 " NA ", " NX ", " NX ";\n\t"                 \
    "mul.rn.f64 " NB ", " NY ", " NY ";\n\t"
#define FRB_STEP7(X, Y, A, B, NX, NY, NA, NB)                \
    "sub.rn.f64 s, " A ", " B ";\n\t"                        \
    "add.rn.f64 t, " X ", " X ";\n\t"                        \
    "mul.rn.f64 t, t, " Y ";\n\t"                            \
    "add.rn.f64 " NX ", s, %10;\n\t"                         \
    "add.rn.f64 " NY ", t, %11;\n\t"                         \
    "mul.rn.f64 " NA ", " NX ", " NX ";\n\t"                 \
    "mul.rn.f64 " NB ", " NY ", " NY ";\n\t"
#define FRB_PAIR_FILTER(A1, B1, A2, B2, TARGET)              \
    "mov.b64 {lo, h1}, " A1 ";\n\t"                          \
    "mov.b64 {lo, h2}, " B1 ";\n\t"                          \
    "mov.b64 {lo, h3}, " A2 ";\n\t"                          \
    "mov.b64 {lo, h4}, " B2 ";\n\t"                          \
    "max.u32 h1, h1, h2;\n\t"                                \
    "max.u32 h3, h3, h4;\n\t"                                \
    "max.u32 h1, h1, h3;\n\t"                                \
    "setp.ge.u32 q, h1, %13;\n\t"                            \
    "@q bra " TARGET ";\n\t"
#define FRB_TILE_LOOP(STEP)                                                                                                 \
    asm volatile("{\n\t"                                                                                                    \
                 ".reg .f64 s, t, x1, y1, a1, b1, x2, y2, a2, b2, x3, y3, a3, b3;\n\t"                                      \
                 ".reg .u32 lo, h1, h2, h3, h4;\n\t"                                                                        \
                 ".reg .pred q;\n\t"                                                                                        \
                 "mov.u32 %9, 4;\n\t"                                                                                       \
                 "setp.ge.u32 q, %8, %12;\n\t"                                                                              \
                 "@q bra FRB_DONE;\n"                                                                                       \
                 "FRB_LOOP:\n\t"                                                                                            \
                 STEP("%0", "%1", "%2", "%3", "x1", "y1", "a1", "b1")                                                       \
                 STEP("x1", "y1", "a1", "b1", "x2", "y2", "a2", "b2")                                                       \
                 FRB_PAIR_FILTER("a1", "b1", "a2", "b2", "FRB_HIT0")                                                        \
                 STEP("x2", "y2", "a2", "b2", "x3", "y3", "a3", "b3")                                                       \
                 STEP("x3", "y3", "a3", "b3", "%0", "%1", "%2", "%3")                                                       \
                 FRB_PAIR_FILTER("a3", "b3", "%2", "%3", "FRB_HIT2")                                                        \
                 "add.u32 %8, %8, 4;\n\t"                                                                                   \
                 "setp.lt.u32 q, %8, %12;\n\t"                                                                              \
                 "@q bra FRB_LOOP;\n\t"                                                                                     \
                 "bra FRB_DONE;\n"                                                                                          \
                 "FRB_HIT0:\n\t"                                                                                            \
                 "mov.f64 %4, x1;\n\tmov.f64 %5, y1;\n\tmov.f64 %6, a1;\n\tmov.f64 %7, b1;\n\t"                             \
                 "mov.f64 %0, x2;\n\tmov.f64 %1, y2;\n\tmov.f64 %2, a2;\n\tmov.f64 %3, b2;\n\t"                             \
                 "mov.u32 %9, 0;\n\t"                                                                                       \
                 "bra FRB_DONE;\n"                                                                                          \
                 "FRB_HIT2:\n\t"                                                                                            \
                 "mov.f64 %4, x3;\n\tmov.f64 %5, y3;\n\tmov.f64 %6, a3;\n\tmov.f64 %7, b3;\n\t"                             \
                 "mov.u32 %9, 2;\n"                                                                                         \
                 "FRB_DONE:\n\t"                                                                                            \
                 "}"                                                                                                        \
                 : "+d"(o.x), "+d"(o.y), "+d"(o.a), "+d"(o.b), "=d"(pre.x), "=d"(pre.y), "=d"(pre.a), "=d"(pre.b), "+r"(i), \
                   "=r"(k)                                                                                                  \
                 : "d"(o.cr), "d"(o.ci), "r"(n4), "r"(filter_hi))
            if (FORM == 6) FRB_TILE_LOOP(FRB_STEP6);
            else FRB_TILE_LOOP(FRB_STEP7);
#undef FRB_TILE_LOOP
#undef FRB_PAIR_FILTER
#undef FRB_STEP7
#undef FRB_STEP6
            if (k == 4u) break;  // i == n4: fewer than four iterations left
            i += k;              // index of the first iteration of the pair
            if (stream_escaped(pre, S)) {
                count = i;
                o.x = pre.x, o.y = pre.y;
                goto escaped;
            }
            ++i;
            if (stream_escaped(o, S)) {
                count = i;
                goto escaped;
            }
            for (++i; (i & 3u) != 0u; ++i) {  // the pair began below n4, so the next multiple of four is at most n4
                stream_core<FORM>(o);
                if (stream_escaped(o, S)) {
                    count = i;
                    goto escaped;
                }
            }
        }
        for (; i < iterations; ++i) {
            stream_core<FORM>(o);
            if (stream_escaped(o, S)) {
                count = i;
                break;
            }
        }
    escaped:;
    }
    // colour straight from the registers; the log2 table (2 KB) is read through the read-only cache: no copy to shared memory and
    // no barrier between a warp's last iteration and its shading
    const uint32_t c = valid ? shade_pixel<true>(p, count, make_double2(o.x, o.y), g_log2_table) : 0u;

    // statistics: loop-body executions (i + 1 | iterations) and the lane slots the lock-step warp spent (32 x its longest orbit).
    // Two same-address atomics per warp (65 000 warps on C1) would serialise in the L2 for longer than the frame takes to
    // iterate, so the warps spread them over 32 pairs of counters that K0 cleared and the host adds up (frb_api.cu: harvest).
    // The kernel only runs caps below 2^16 (make_plan), so a warp's sum fits a word.
    {
        const uint32_t useful = (valid && p.iterations != 0u) ? min(count, p.iterations - 1u) + 1u : 0u;
        const uint32_t longest = __reduce_max_sync(0xffffffffu, useful);
        const uint32_t sum = __reduce_add_sync(0xffffffffu, useful);
        if (lane == 0 && sum != 0u) {
            unsigned long long *part = p.stats + kTileStatsWord + 2u * ((blockIdx.x + 5u * blockIdx.y + (threadIdx.x >> 5)) & (kTileStatsPairs - 1u));
            atomicAdd(part, (unsigned long long)sum);
            atomicAdd(part + 1, 32ull * longest);
        }
    }

    if (p.format == 4u) {
        if (valid) reinterpret_cast<uint32_t *>(p.out)[(size_t)lr * p.width + x] = c | 0xff000000u;
        return;
    }
    // RGB24: a tile row is 3 TW bytes = 3 TW / 4 words.  Word j of the row takes its bytes from pixels p0 = 4j / 3 and p0 + 1
    // (two shuffles and a funnel shift), and lane ty * TW + j stores it: consecutive lanes, consecutive words.  Needs the row
    // segment to start on a word boundary: width a multiple of 4 (x0 is one of TW) and an aligned image.
    const bool valid_row = lr < p.rows;
    if (x0 + TW <= p.width && (p.width & 3u) == 0u && (reinterpret_cast<uintptr_t>(p.out) & 3u) == 0u) {
        const uint32_t p0 = (tx * 4u) / 3u, sh = (tx * 4u - p0 * 3u) * 8u;
        const uint32_t c0 = __shfl_sync(0xffffffffu, c, (int)(ty * TW + p0));
        const uint32_t c1 = __shfl_sync(0xffffffffu, c, (int)(ty * TW + min(p0 + 1u, (uint32_t)TW - 1u)));
        const uint32_t word = __funnelshift_r(c0 | (c1 << 24), c1 >> 8, sh);
        if (tx < (3u * TW) / 4u && valid_row)
            reinterpret_cast<uint32_t *>(p.out + ((size_t)lr * p.width + x0) * 3u)[tx] = word;
        return;
    }
    if (valid) {
        uint8_t *q = p.out + ((size_t)lr * p.width + x) * 3u;
        q[0] = (uint8_t)c;
        q[1] = (uint8_t)(c >> 8);
        q[2] = (uint8_t)(c >> 16);
    }
}

}  // namespace frb
