// frb_kernels.cuh — device side of the B200 escape-time renderer (sm_100a only).
//
// The path being replaced (reference, /root/reference):
//   calc/src/lib.rs:181-197  coord_to_space / xy_to_imaginary      -> frb_coord_tables_kernel
//   calc/src/lib.rs:245-257  recursive (the z <- z^2 + c loop)     -> frb_iterate_stream_kernel / frb_iterate_simple_kernel
//   calc/src/lib.rs:214-234  classify + smooth colour + palette    -> frb_shade_kernel
//   calc/src/lib.rs:133-139  color_multiply (with the g/b swap)    -> shade_pixel()
//   src/lib.rs:255-269       get_image row loop (rayon)            -> the persistent chunk queue
//
// All FP64 arithmetic on the result path uses the round-to-nearest intrinsics (__dadd_rn, __dmul_rn, __fma_rn,
// __ddiv_rn, __dsqrt_rn), which nvcc never contracts or reassociates; the translation unit is additionally built
// with -fmad=false.  Tensor cores and TMA are deliberately unused: the path is a scalar FP64 recurrence with a
// data-dependent exit, bounded by the FP64 pipe (64 lanes / SM), not by HBM or by a contraction.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace frb {

constexpr int kIterateThreads = 256;  // 8 warps per CTA; warps are fully independent (no smem, no CTA barrier)
constexpr int kUnroll = 8;            // iterations per lock-step block between refill checks
constexpr int kShadeThreads = 256;
constexpr int kShadePxPerThread = 4;  // 4 px/thread -> one 16-byte RGBA store, 3 x 16-byte RGB stores per 4 threads

enum : uint32_t { kAlgoMandelbrot = 0, kAlgoFern = 1, kAlgoJulia = 2 };

// Everything a launch needs, passed by value (__grid_constant__ -> constant bank, uniform loads).
struct RenderParams {
    // geometry of this launch: `rows` compact local rows of a shard, each `width` pixels
    uint32_t width;
    uint32_t height;
    uint32_t rows;
    uint32_t band_rows;    // >= 1
    uint32_t shard_index;
    uint32_t shard_count;  // >= 1
    uint32_t iterations;
    uint32_t algo;
    // escape test
    double limit_sq;       // S = limit * limit (calc/src/lib.rs:246)
    uint32_t filter_hi;    // high word of S/2: max(hi(re^2), hi(im^2)) >= filter_hi is necessary for re^2+im^2 > S
    uint32_t chunk;        // pixels per queue chunk (a segment of one row)
    uint32_t chunks_per_row;
    uint32_t total_chunks;
    double julia_re, julia_im;
    // coordinate map inputs (calc/src/lib.rs:181-197)
    double pos_re, pos_im, scale_re, scale_im;
    // colour (calc/src/lib.rs:214-234)
    double stable_limit, exposure;
    uint32_t smooth, inside;
    uint32_t primary;      // stored fields r | g<<8 | b<<16
    uint32_t secondary;
    uint32_t format;       // 3 = RGB24, 4 = RGBA32
    uint32_t filter_abs_hi; // high word of sqrt(S/2), minus 1: max(|re|,|im|) >= this is necessary for re^2+im^2 > S/2
    uint32_t chunk_begin;   // first chunk of this launch (row slabs of one shard are separate launches); 0 normally
    uint32_t defer_block;   // K1c/K1r: iterations per block of the deferred kernel (a record is at most this + 1 steps short)
    uint32_t resolve_row0;  // K1r: first local row and number of rows to resolve
    uint32_t resolve_rows;
    uint32_t records;       // K2: 1 = iters/zfinal hold K1c's provisional records; resolve them while shading (local row 0 of the
                            // pointers is shard row resolve_row0)
    uint32_t stream_gen;    // K0 -> K2s: generation of this frame, published once the control words are reset (0 = not streamed)
    // K1c's queue (frb_deferred.cuh): the launch covers local rows [q_row0, ...); the first q_nbig chunks are segments of `chunk`
    // pixels (chunks_per_row per row, q_rows_big rows), the rest — the END PHASE of the frame — segments of q_chunk2 pixels
    // (q_cpr2 per row).  total_chunks counts both.  q_pool_max: a drained warp with at most this many live lanes hands them to
    // the CTA's orbit pool and leaves (0 = no pool).
    uint32_t q_row0, q_rows_big, q_nbig, q_cpr2, q_chunk2, q_pool_max;
    uint32_t band_shift;    // log2(band_rows) when band_rows is a power of two (the default bands of 16 rows), else 32
    uint32_t pad_band;
    // Output addressing of K2 / K2s.  out_band_q == 0: `out` is this launch's own compact image.  Otherwise `out` is where the
    // shard's band 0 starts inside a WHOLE frame shared with other devices' shards (frb_render, device = -1): compact byte offset
    // o of the shard maps to o + (o / band_bytes) * out_band_gap, band_bytes = 16 * out_band_q (one band of kDefaultBandRows rows),
    // and out_off is the compact offset of this launch's first pixel (row slabs are launches of their own).
    uint32_t out_band_q;    // band_bytes / 16 (bands are multiples of 16 bytes: 16 rows)
    unsigned long long out_band_gap;  // bytes between the end of a band and the shard's next band = (G - 1) * band_bytes
    unsigned long long out_off;
    // colour fast path (shade_pixel): uniform quantities converted once on the host instead of once per pixel
    uint32_t fast_shade;    // 1 = exposure is finite and moderate, iterations >= 1: products of the fast path stay finite
    double iterations_d;    // (double)iterations
    double inv_iterations;  // RN(1 / iterations_d)
    double prim_r, prim_g, prim_b;  // STORED fields of primary_color as doubles (the g/b swap happens in the multiply)
    // buffers
    double *tab_re;              // [width]   re0(x)
    double *tab_im;              // [height]  im0(y), indexed by IMAGE row
    uint32_t *iters;             // [rows*width] count returned by recursive()
    double2 *zfinal;             // [rows*width] position returned by recursive()
    uint8_t *out;                // [rows*width*format]
    unsigned int *queue;         // chunk counter
    uint32_t *wm;                // K1c: per-warp watermarks for the streaming finisher (frb_stream.cuh); nullptr = not published
    unsigned long long *stats;   // [0] useful iterations, [1] lane-slot iterations issued (32 x warp trips x unroll)
};

// Address of compact byte offset `off` of this launch's image (see RenderParams::out_band_q).  Stores of up to 16 bytes at
// 16-byte-aligned offsets never straddle a band: bands are multiples of 16 bytes.
__device__ __forceinline__ uint8_t *out_ptr(const RenderParams &p, unsigned long long off)
{
    if (p.out_band_q == 0u) return p.out + off;
    const unsigned long long o = p.out_off + off;
    const uint32_t band = (uint32_t)(o >> 4) / p.out_band_q;  // o < 2^36 (2^32 pixels x 4 B, frb_api.cu: make_plan)
    return p.out + o + (unsigned long long)band * p.out_band_gap;
}

// The launch parameters of the sub-image starting at compact pixel pix0 (row slabs): its output begins there.
__host__ __device__ inline void offset_out(RenderParams &sub, const RenderParams &p, size_t pix0)
{
    if (p.out_band_q != 0u) sub.out_off = p.out_off + (unsigned long long)pix0 * p.format;
    else sub.out = p.out + pix0 * p.format;
}

// local (compact) row of a shard -> image row; see FrbShard in include/frb.h
__device__ __forceinline__ uint32_t local_row_to_y(const RenderParams &p, uint32_t lr)
{
    if (p.shard_count == 1u) return lr;  // one shard: its rows are the image's
    if (p.band_shift < 32u) {
        const uint32_t band = lr >> p.band_shift;
        return ((band * p.shard_count + p.shard_index) << p.band_shift) + (lr & (p.band_rows - 1u));
    }
    const uint32_t band = lr / p.band_rows;
    const uint32_t r = lr - band * p.band_rows;
    return (band * p.shard_count + p.shard_index) * p.band_rows + r;
}

// ---------------------------------------------------------------------------------------------------------
// K0: coordinate tables.  re0 depends on x only and im0 on y only (calc/src/lib.rs:194-195), so W+H values
// computed once with the reference's exact operation order replace 4 divisions per pixel.  Also resets the
// work queue and the statistics (stream-ordered before the iterate kernel).
// ---------------------------------------------------------------------------------------------------------
__global__ void frb_coord_tables_kernel(const __grid_constant__ RenderParams p)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < 16) p.queue[t] = 0u;  // one counter per row slab (frb_api.cu: enqueue_sliced)
    if (t < 64) p.stats[16 + t] = 0ull;  // the tile kernel's spread counters (frb_tile.cuh: kTileStatsWord, kTileStatsPairs)
    if (t == 0) {
        p.stats[0] = 0ull;
        p.stats[1] = 0ull;
        p.stats[2] = 0ull;  // streaming finisher: status (low word), published minimum watermark (high word)
        p.stats[10] = 0ull; //                     watcher election (low word), next work item (high word)
        p.stats[12] = 0ull; //                     finisher CTAs that saw the frame start
        if (p.stream_gen != 0u) {
            // the finisher of this frame is already resident (launched first, so that its CTAs are the older ones) and sleeps on
            // this word; the watermark table was cleared by the memset that precedes this kernel in stream order
            __threadfence();
            *reinterpret_cast<volatile uint32_t *>(p.stats + 11) = p.stream_gen;
        }
#ifdef FRB_STREAM_TRACE
        p.stats[3] = ~0ull; p.stats[4] = ~0ull; p.stats[5] = 0ull; p.stats[6] = 0ull; p.stats[7] = ~0ull;
        p.stats[9] = ~0ull; p.stats[13] = ~0ull;  // queue drained (first warp to see it), first iterate warp running
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(p.stats[8]));
#endif
    }
    const double wf = (double)p.width, hf = (double)p.height;
    if (t < p.width) {
        // coord_to_space(x, height, (width/height)/2.0, pos.re, scale.re)
        const double off = __ddiv_rn(__ddiv_rn(wf, hf), 2.0);
        const double v = __ddiv_rn(__dsub_rn(__ddiv_rn((double)t, hf), off), p.scale_re);
        p.tab_re[t] = __dadd_rn(v, p.pos_re);
    }
    if (t < p.height) {
        // coord_to_space(y, height, 0.5, pos.im, scale.im)
        const double v = __ddiv_rn(__dsub_rn(__ddiv_rn((double)t, hf), 0.5), p.scale_im);
        p.tab_im[t] = __dadd_rn(v, p.pos_im);
    }
}

// ---------------------------------------------------------------------------------------------------------
// One iteration, two formulations.
// ---------------------------------------------------------------------------------------------------------

// Per-lane orbit state.  (x, y) = z_k; (a, b) = (x*x, y*y) carried from the previous escape test (stream kernels).
struct Orbit {
    double x, y, a, b, cr, ci;
};

// Reference order exactly as written (calc/src/lib.rs:88-89, 103-104, 95, 251): 6 mul + 4 add + 1 compare.
// Returns true when `next.squared_distance() > squared`; (x, y) is `next` afterwards either way.
__device__ __forceinline__ bool step_exact(Orbit &o, const double S)
{
    const double a = __dmul_rn(o.x, o.x);
    const double b = __dmul_rn(o.y, o.y);
    double nre = __dsub_rn(a, b);
    const double t = __dmul_rn(2.0, o.x);
    double nim = __dmul_rn(t, o.y);
    nre = __dadd_rn(nre, o.cr);
    nim = __dadd_rn(nim, o.ci);
    o.x = nre;
    o.y = nim;
    const double d = __dadd_rn(__dmul_rn(nre, nre), __dmul_rn(nim, nim));
    return d > S;
}

// The streaming formulation: the two squares of squared_distance() are carried into the next iteration's
// square() (they are the same roundings of the same operands), and the sum d = a' + b' with its compare is only
// evaluated when an integer-pipe filter fires:
//   max(hi32(a'), hi32(b')) >= hi32(S/2)   is NECESSARY for   a' + b' > S
// (a', b' are squares, so non-negative or NaN; if both are < S/2 the exact sum is < S and rounding is monotone;
// the unsigned compare of the high words also fires for +inf and NaN, and the exact test then decides exactly as
// the reference would).  That leaves per iteration:
//   s   = a - b            (re*re) - (im*im)                                   DADD
//   nre = s + cr           + c.re                                               DADD
//   a'  = nre * nre ; b' = nim * nim                                            2 x DMUL
// and for the imaginary part either
//   FORM 7 (exact for ANY input):  t = x + x (== 2.0*x, exact) ; q = t * y ; nim = q + ci        DADD DMUL DADD
//   FORM 6:  p = x * y ; nim = fma(p, 2.0, ci)                                                   DMUL DFMA
//     2*p is exact, so the fma rounds 2*round(x*y) + ci once; 2*round(x*y) == round((2x)*y) unless x*y is
//     subnormal or overflows.  The host (frb_api.cu: form6_is_exact) only selects FORM 6 when neither can change a
//     result bit: S <= 1e300 and |c.im| is either 0 with im(z) == 0 forever (Mandelbrot real-axis row) or
//     >= 2^-900 (then a subnormal x*y is far below half an ulp of c.im and both forms round to the same value).
// 6 is the floor for bit-exactness: the reference performs six roundings per iteration whose results are all
// live (a-b, +c.re, 2xy, +c.im, re^2, im^2) and every FP64 instruction delivers exactly one.
template <int FORM>
__device__ __forceinline__ void stream_core(Orbit &o)
{
    const double s = __dsub_rn(o.a, o.b);
    double nim;
    if (FORM == 6) {
        const double p = __dmul_rn(o.x, o.y);
        nim = __fma_rn(p, 2.0, o.ci);
    } else {
        const double t = __dadd_rn(o.x, o.x);
        nim = __dadd_rn(__dmul_rn(t, o.y), o.ci);
    }
    const double nre = __dadd_rn(s, o.cr);
    o.x = nre;
    o.y = nim;
    o.a = __dmul_rn(nre, nre);
    o.b = __dmul_rn(nim, nim);
}

// integer-pipe escape filter (necessary condition, see above)
__device__ __forceinline__ bool stream_filter(const Orbit &o, const uint32_t filter_hi)
{
    const uint32_t ha = (uint32_t)__double2hiint(o.a);
    const uint32_t hb = (uint32_t)__double2hiint(o.b);
    return max(ha, hb) >= filter_hi;
}

// the reference's test, `next.squared_distance() > squared` (calc/src/lib.rs:250-251), on the carried squares
__device__ __forceinline__ bool stream_escaped(const Orbit &o, const double S)
{
    return __dadd_rn(o.a, o.b) > S;
}

// ---------------------------------------------------------------------------------------------------------
// K1: persistent lane-refill iterate kernel.
//
// Grid = SMs x resident CTAs; every WARP is an independent worker.  Work units are chunks (segments of one
// image row, p.chunk pixels) handed out by one global atomic counter; the id of the next chunk is prefetched
// so the atomic's latency is hidden behind iteration blocks.  Each lane owns one pixel's orbit in registers.
// The warp runs kUnroll iterations in lock step, then (warp-uniform ballot) lanes whose pixel finished store
// its raw result (count, final z: 20 B) and take the next pixel of the warp's chunk cursor.  Lanes therefore
// never wait for the slowest pixel of a tile: utilisation of the FP64 pipe is bounded by the U/2 iterations a
// finished lane idles until the next refill point, not by the spread of escape times inside a warp.
// Colour is NOT computed here (its sqrt/log2 would run at 1/32 lane efficiency); frb_shade_kernel does it at full
// efficiency from the raw results, which also gives the parity hook (frb_render_counts) for free.
// ---------------------------------------------------------------------------------------------------------
template <int FORM>
__global__ void __launch_bounds__(kIterateThreads, 4)
frb_iterate_stream_kernel(const __grid_constant__ RenderParams p)
{
    constexpr unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;

    const double S = p.limit_sq;
    const uint32_t filter_hi = p.filter_hi;
    const uint32_t iterations = p.iterations;
    const bool julia = (p.algo == kAlgoJulia);

    Orbit o = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    uint32_t i = 0;       // iterations already executed for the lane's pixel
    uint32_t pix = 0;     // compact pixel index (output slot)
    bool active = false;

    // warp-uniform chunk cursor
    uint32_t cur_x = 0, cur_end = 0, cur_rowpix = 0;
    double cur_im = 0.0;
    bool drained = false;
    // prefetched id of the next chunk (meaningful on lane 0)
    uint32_t next_chunk = 0;
    if (lane == 0) next_chunk = atomicAdd(p.queue, 1u);

    uint32_t trips = 0;             // lock-step blocks issued by this warp

    for (;;) {
        unsigned need = __ballot_sync(FULL, !active);
        if (need != 0u) {
            while (!drained && need != 0u) {
                if (cur_x == cur_end) {
                    const uint32_t c = __shfl_sync(FULL, next_chunk, 0);
                    if (c >= p.total_chunks) {
                        drained = true;
                        break;
                    }
                    if (lane == 0) next_chunk = atomicAdd(p.queue, 1u);
                    const uint32_t cg = c + p.chunk_begin;
                    const uint32_t lr = cg / p.chunks_per_row;
                    const uint32_t seg = cg - lr * p.chunks_per_row;
                    cur_x = seg * p.chunk;
                    cur_end = min(cur_x + p.chunk, p.width);
                    cur_rowpix = lr * p.width;
                    cur_im = __ldg(p.tab_im + local_row_to_y(p, lr));
                }
                const uint32_t avail = cur_end - cur_x;
                const uint32_t rank = __popc(need & lt_mask);
                if (!active && rank < avail) {
                    const uint32_t px = cur_x + rank;
                    const double re0 = __ldg(p.tab_re + px);
                    // get_recursive_pixel (calc/src/lib.rs:208-210): Mandelbrot c = start, Julia c = julia_set
                    o.x = re0;
                    o.y = cur_im;
                    o.cr = julia ? p.julia_re : re0;
                    o.ci = julia ? p.julia_im : cur_im;
                    o.a = __dmul_rn(re0, re0);
                    o.b = __dmul_rn(cur_im, cur_im);
                    i = 0;
                    pix = cur_rowpix + px;
                    active = true;
                }
                cur_x += min((uint32_t)__popc(need), avail);
                need = __ballot_sync(FULL, !active);
            }
            if (need == FULL) break;  // queue drained and no lane holds a pixel
        }

        ++trips;
        if (active) {
            const uint32_t rem = iterations - i;
            bool done = false;
            uint32_t count = iterations;
            if (rem >= (uint32_t)kUnroll) {
                // hot path: kUnroll x (6|7 FP64 + filter + branch); the exact test lives on the cold path only
                uint32_t k = 0;
                bool cand = false;
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    stream_core<FORM>(o);
                    if (stream_filter(o, filter_hi)) {
                        k = (uint32_t)u;
                        cand = true;
                        break;
                    }
                }
                if (cand) {
                    // cold path: finish this block one step at a time with the exact test after every step
                    for (;;) {
                        if (stream_escaped(o, S)) {
                            count = i + k;
                            done = true;
                            break;
                        }
                        if (++k == (uint32_t)kUnroll) break;
                        stream_core<FORM>(o);
                    }
                }
                i += (uint32_t)kUnroll;
            } else {
                // fewer than kUnroll iterations left: finish the pixel here
                done = true;
                for (uint32_t k = 0; k < rem; ++k) {
                    stream_core<FORM>(o);
                    if (stream_filter(o, filter_hi) && stream_escaped(o, S)) {
                        count = i + k;
                        break;
                    }
                }
            }
            if (done) {
                // recursive() returns (next, i) on escape, (previous, iterations) otherwise: (o.x, o.y) is that position
                p.iters[pix] = count;
                p.zfinal[pix] = make_double2(o.x, o.y);
                active = false;
            }
        }
    }

    // statistics: one atomic per warp (the useful-iteration total is summed by the shade / sum kernel)
    if (lane == 0) atomicAdd(p.stats + 1, (unsigned long long)trips * (unsigned long long)(32 * kUnroll));
}

}  // namespace frb

#include "frb_deferred.cuh"  // K1c + K1r: the default — no recorded positions, escapes resolved by a second pass

namespace frb {

// ---------------------------------------------------------------------------------------------------------
// K1s: cross-check kernel — one thread per pixel, reference order as written, no queue, no refill.
// Lock-step warps of 32x1 pixels, so slower than the stream kernels; used to validate the stream kernels at full
// frame sizes where the CPU oracle would need minutes, and available as FRB_KERNEL_SIMPLE.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
frb_iterate_simple_kernel(const __grid_constant__ RenderParams p)
{
    // blockIdx.x = local row (x dimension allows 2^31-1 rows), blockIdx.y = column block
    const uint32_t x = blockIdx.y * blockDim.x + threadIdx.x;
    const uint32_t lr = blockIdx.x;
    const bool valid = (x < p.width) && (lr < p.rows);
    if (valid) {
        const double re0 = p.tab_re[x];
        const double im0 = p.tab_im[local_row_to_y(p, lr)];
        const bool julia = (p.algo == kAlgoJulia);
        Orbit o = {re0, im0, 0.0, 0.0, julia ? p.julia_re : re0, julia ? p.julia_im : im0};
        const double S = p.limit_sq;
        uint32_t count = p.iterations;
        for (uint32_t i = 0; i < p.iterations; ++i) {
            if (step_exact(o, S)) {
                count = i;
                break;
            }
        }
        const uint32_t pix = lr * p.width + x;
        p.iters[pix] = count;
        p.zfinal[pix] = make_double2(o.x, o.y);
    }
}

// ---------------------------------------------------------------------------------------------------------
// K2: shade.  Raw (count, final z) -> colour, at full lane efficiency, 16-byte coalesced stores.
// ---------------------------------------------------------------------------------------------------------

// Rust `f64 as u8`: truncate toward zero, saturate to [0, 255], NaN -> 0.  cvt.rzi.u32.f64 saturates negatives to 0
// and large values to 2^32-1; NaN is handled explicitly (measured on B200: the conversion does not return 0 for it).
__device__ __forceinline__ uint32_t sat_u8(double v)
{
    return (v == v) ? min(__double2uint_rz(v), 255u) : 0u;
}

// color_multiply (calc/src/lib.rs:133-139): RGB::new(r*m, g*m, b*m) with new(r, b, g) => out.g = b*m, out.b = g*m.
__device__ __forceinline__ uint32_t color_multiply(uint32_t stored, double m)
{
    const double cr = (double)(stored & 0xffu);
    const double cg = (double)((stored >> 8) & 0xffu);
    const double cb = (double)((stored >> 16) & 0xffu);
    const uint32_t out_r = sat_u8(__dmul_rn(cr, m));
    const uint32_t out_b = sat_u8(__dmul_rn(cg, m));  // second ctor argument lands in .b
    const uint32_t out_g = sat_u8(__dmul_rn(cb, m));  // third ctor argument lands in .g
    return out_r | (out_g << 8) | (out_b << 16);
}

// ---- log2 for the smooth-colour term ----------------------------------------------------------------------------------
// libdevice's log2(double) costs ~45 FP64-pipe instructions (denormal / special handling, a division by Newton steps,
// compensated reconstruction); smooth colouring calls it twice per pixel plus a square root and two divisions — ~200
// FP64 instructions a pixel, 5x the ITERATIONS of a C1 pixel.  The arguments here are tame (dist >= 2 and finite,
// log_zn in [0.25, 256)), so: split x = 2^e * m, look up {1/c, log2 c} for the 1/128-wide interval m falls in, and
// log2 x = e + log2 c + log2(1 + r),  r = fma(m, 1/c, -1),  |r| < 2^-8, Taylor to degree 6 (truncation < 2e-18).
// Error <= 2.4e-16 * max(1, |log2 x|) (scripts/gen_log2_table.py builds the table; checked against long double on 4M points),
// i.e. the same last-bit noise libdevice has against glibc/Rust — colour stays within the +-1 LSB the boundary allows, and
// tests/test_parity_gpu.py::test_fast_shade_matches_libdevice_shade counts how rarely a byte differs at all.
__device__ const double2 g_log2_table[128] = {
#include "frb_log2_table.inc"
};

// degree-6 Taylor coefficients of log2(1 + r) from r^6 down to r^1, then 2^52 + 1023 (the exponent bias of the 2^52 trick)
__constant__ double c_log2_poly[7] = {-0.24044917348149392, 0.28853900817779266, -0.36067376022224085, 0.48089834696298783,
                                      -0.7213475204444817,  1.4426950408889634,  4503599627371519.0};

__device__ __forceinline__ void load_log2_table(double2 *dst)  // dst: 128 entries of shared memory; caller syncs
{
    for (unsigned i = threadIdx.x; i < 128u; i += blockDim.x) dst[i] = g_log2_table[i];
}

// x positive, normal, finite.  GLOBAL_TABLE: `tbl` is g_log2_table itself, read through the read-only cache (the tile kernel);
// otherwise a copy in shared memory.
template <bool GLOBAL_TABLE = false>
__device__ __forceinline__ double shade_log2(double x, const double2 *tbl)
{
    const uint32_t hi = (uint32_t)__double2hiint(x);
    const double m = __hiloint2double((int)((hi & 0x000fffffu) | 0x3ff00000u), __double2loint(x));
    const double2 t = GLOBAL_TABLE ? __ldg(tbl + ((hi >> 13) & 127u)) : tbl[(hi >> 13) & 127u];
    const double r = fma(m, t.x, -1.0);
    // biased exponent -> double without a conversion instruction: the bits of 2^52 + E, minus (2^52 + 1023)
    // the coefficients come from the constant bank as instruction operands; as literals each cost two UMOVs per use site
    const double e = __dsub_rn(__hiloint2double(0x43300000, (int)(hi >> 20)), c_log2_poly[6]);
    double q = fma(r, c_log2_poly[0], c_log2_poly[1]);
    q = fma(r, q, c_log2_poly[2]);
    q = fma(r, q, c_log2_poly[3]);
    q = fma(r, q, c_log2_poly[4]);
    q = fma(r, q, c_log2_poly[5]);
    return fma(r, q, __dadd_rn(e, t.y));
}

// get_recursive_pixel after recursive() (calc/src/lib.rs:214-234).  Returns r | g<<8 | b<<16.  `lg` = the log2 table in
// shared memory.
template <bool GLOBAL_TABLE = false>
__device__ __forceinline__ uint32_t shade_pixel(const RenderParams &p, uint32_t count, double2 z, const double2 *lg)
{
    if (p.algo == kAlgoFern) return 0u;  // `_ => return RGB::BLACK`
    const double dist = __dadd_rn(__dmul_rn(z.x, z.x), __dmul_rn(z.y, z.y));
    if (dist > p.stable_limit) {
        double it = (double)count;
        bool fast = p.fast_shade != 0u;
        if (p.smooth) {
            // 2 <= dist < inf  (high words 0x40000000 .. 0x7fefffff): both logarithms take the table path
            if (fast && (uint32_t)__double2hiint(dist) - 0x40000000u < 0x3ff00000u) {
                const double log_zn = __dmul_rn(shade_log2<GLOBAL_TABLE>(dist, lg), 0.25);  // log2(sqrt(dist)) / 2
                it = __dadd_rn(it, __dsub_rn(1.0, shade_log2<GLOBAL_TABLE>(log_zn, lg)));
            } else {
                const double log_zn = __ddiv_rn(log2(__dsqrt_rn(dist)), 2.0);
                const double nu = log2(log_zn);
                it = __dadd_rn(it, __dsub_rn(1.0, nu));
                fast = false;  // `it` may be infinite or NaN here
            }
        }
        if (fast) {
            // it / iterations, correctly rounded without the division sequence: q within 1 ulp, exact remainder, one
            // correction (Markstein; needs RN(1/N), finite operands and a divisor whose mantissa is not all ones — an
            // integer below 2^32).  Everything is finite, so the NaN checks of sat_u8 are not needed either;
            // cvt.rzi.u32.f64 saturates negatives to 0.
            double q = __dmul_rn(it, p.inv_iterations);
            q = fma(fma(-q, p.iterations_d, it), p.inv_iterations, q);
            const double mult = __dmul_rn(q, p.exposure);
            const uint32_t out_r = min(__double2uint_rz(__dmul_rn(p.prim_r, mult)), 255u);
            const uint32_t out_b = min(__double2uint_rz(__dmul_rn(p.prim_g, mult)), 255u);  // color_multiply's g/b swap
            const uint32_t out_g = min(__double2uint_rz(__dmul_rn(p.prim_b, mult)), 255u);
            return out_r | (out_g << 8) | (out_b << 16);
        }
        const double mult = __dmul_rn(__ddiv_rn(it, (double)p.iterations), p.exposure);
        return color_multiply(p.primary, mult);
    }
    if (p.inside) return color_multiply(p.secondary, dist);
    return 0u;
}

// Shades the 1024-pixel blocks first, first + stride, ... of the p.rows x p.width pixels the pointers of `p` address
// (one CTA of kShadeThreads threads; `stage` = kShadeThreads * 3 words of shared memory, `lg` = the log2 table).  Adds the loop-body executions of
// calc::recursive (i + 1 for an escape at index i, else iterations) of the pixels it shaded to `useful`.
__device__ __forceinline__ void shade_blocks(const RenderParams &p, uint32_t *stage, const double2 *lg, uint64_t first,
                                             uint64_t stride, unsigned long long &useful)
{
    const uint64_t npix = (uint64_t)p.rows * p.width;
    const uint64_t cta_px = (uint64_t)kShadeThreads * kShadePxPerThread;
    const uint64_t nblocks = (npix + cta_px - 1) / cta_px;
    for (uint64_t blk = first; blk < nblocks; blk += stride) {
        const uint64_t base = blk * cta_px + (uint64_t)threadIdx.x * kShadePxPerThread;
        uint32_t c[4] = {0u, 0u, 0u, 0u};
        // position of the thread's first pixel (only needed to look up c when provisional records are resolved here)
        uint32_t lr = 0, px = 0;
        double c_im = 0.0;
        const bool lookup = p.records != 0u && p.algo != kAlgoJulia && base < npix;
        if (lookup) {
            const uint32_t row = (uint32_t)(base / p.width);
            px = (uint32_t)(base - (uint64_t)row * p.width);
            lr = p.resolve_row0 + row;
            c_im = __ldg(p.tab_im + local_row_to_y(p, lr));
        }
        if (base + 4 <= npix) {
            const uint4 n4 = *reinterpret_cast<const uint4 *>(p.iters + base);
            uint32_t n[4] = {n4.x, n4.y, n4.z, n4.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                double2 z = p.zfinal[base + k];
                if (p.records != 0u) {
                    double cr = p.julia_re, ci = p.julia_im;
                    if (lookup) {
                        cr = __ldg(p.tab_re + px);
                        ci = c_im;
                        if (++px == p.width && k < 3) {
                            px = 0;
                            c_im = __ldg(p.tab_im + local_row_to_y(p, ++lr));
                        }
                    }
                    resolve_record(p, n[k], z, cr, ci);
                }
                c[k] = shade_pixel(p, n[k], z, lg);
                useful += min(n[k], p.iterations - 1u) + 1u;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (base + k < npix) {
                    uint32_t n = p.iters[base + k];
                    double2 z = p.zfinal[base + k];
                    if (p.records != 0u) {
                        double cr = p.julia_re, ci = p.julia_im;
                        if (lookup) {
                            cr = __ldg(p.tab_re + px);
                            ci = c_im;
                            if (++px == p.width && base + k + 1 < npix) {
                                px = 0;
                                c_im = __ldg(p.tab_im + local_row_to_y(p, ++lr));
                            }
                        }
                        resolve_record(p, n, z, cr, ci);
                    }
                    c[k] = shade_pixel(p, n, z, lg);
                    useful += min(n, p.iterations - 1u) + 1u;
                }
        }
        const bool full_cta = (blk + 1) * cta_px <= npix;
        if (p.format == 4u) {
            if (base + 4 <= npix) {
                uint4 v = make_uint4(c[0] | 0xff000000u, c[1] | 0xff000000u, c[2] | 0xff000000u, c[3] | 0xff000000u);
                *reinterpret_cast<uint4 *>(out_ptr(p, base * 4)) = v;
            } else {
                for (int k = 0; k < 4; ++k)
                    if (base + k < npix) *reinterpret_cast<uint32_t *>(out_ptr(p, (base + k) * 4)) = c[k] | 0xff000000u;
            }
        } else {
            // pack 4 px (12 B) as three words: r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
            const uint32_t w0 = c[0] | (c[1] << 24);
            const uint32_t w1 = (c[1] >> 8) | (c[2] << 16);
            const uint32_t w2 = (c[2] >> 16) | (c[3] << 8);
            if (full_cta) {
                stage[threadIdx.x * 3 + 0] = w0;
                stage[threadIdx.x * 3 + 1] = w1;
                stage[threadIdx.x * 3 + 2] = w2;
                __syncthreads();
                // 256 threads x 12 B = 3072 B = 192 x 16 B
                if (threadIdx.x < (kShadeThreads * 3) / 4)
                    *reinterpret_cast<uint4 *>(out_ptr(p, blk * cta_px * 3 + threadIdx.x * 16u)) = reinterpret_cast<const uint4 *>(stage)[threadIdx.x];
                __syncthreads();
            } else {
                const uint32_t w[3] = {w0, w1, w2};
                for (int k = 0; k < 12; ++k)
                    if (base * 3 + k < npix * 3) *out_ptr(p, base * 3 + k) = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
            }
        }
    }
}

// One warp-reduced atomic per warp into stats[0] (the G iter/s numerator).
__device__ __forceinline__ void shade_publish_useful(const RenderParams &p, unsigned long long useful)
{
    if (p.iterations == 0u || p.algo == kAlgoFern) useful = 0;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) useful += __shfl_xor_sync(0xffffffffu, useful, s);
    if ((threadIdx.x & 31u) == 0 && useful != 0ull) atomicAdd(p.stats + 0, useful);
}

__global__ void __launch_bounds__(kShadeThreads)
frb_shade_kernel(const __grid_constant__ RenderParams p)
{
    __shared__ __align__(16) uint32_t stage[kShadeThreads * 3];  // 12 B per thread for RGB24
    __shared__ double2 lg[128];
    load_log2_table(lg);
    __syncthreads();
    unsigned long long useful = 0;
    shade_blocks(p, stage, lg, blockIdx.x, gridDim.x, useful);
    shade_publish_useful(p, useful);
}

// Sum of loop-body executions when no image is shaded (frb_render_counts).
__global__ void __launch_bounds__(256)
frb_sum_iterations_kernel(const __grid_constant__ RenderParams p)
{
    const uint64_t npix = (uint64_t)p.rows * p.width;
    unsigned long long useful = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (uint64_t)gridDim.x * blockDim.x)
        useful += min(p.iters[i], p.iterations - 1u) + 1u;
    if (p.iterations == 0u || p.algo == kAlgoFern) useful = 0;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) useful += __shfl_xor_sync(0xffffffffu, useful, s);
    if ((threadIdx.x & 31u) == 0 && useful != 0ull) atomicAdd(p.stats + 0, useful);
}

}  // namespace frb

#include "frb_stream.cuh"  // K2s: shading + copy-out streamed beside K1c
#include "frb_tile.cuh"    // K1t: short iteration caps — lock-step tiles, colour in the same kernel
