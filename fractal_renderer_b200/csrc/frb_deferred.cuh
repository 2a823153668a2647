// frb_deferred.cuh — K1c: persistent lane-refill iterate kernel with DEFERRED escape resolution, and K1r, the
// resolve pass that follows it.  Included by frb_kernels.cuh.
//
// Measured on B200 (profiles/r1_sparse6_v4_ncu_summary.txt): an SMSP issues one instruction per cycle, an FP64
// instruction holds the issue port for two, and the v4 sparse kernel spends 11.9 % of all issue cycles on non-FP64
// instructions — half in the loop (5 per block of 8 iterations), half in the per-pixel service code, most of which is
// the search through the 8 recorded positions of a lane's last block (28 selects) and the tail loop.  This kernel
// removes the recorded positions altogether:
//
//   block = B x { DADD s=a-b ; DMUL q=x*y ; DADD x'=s+c.re ; DFMA y'=fma(q,2,c.im) ; DMUL a=x'^2 ; DMUL b=y'^2 }
//           IADD cnt+=B ; VIMNMX3 m=max(hi(a),hi(b),cnt) ; ISETP m>=filter ; @P BRA stop_k          (4 non-FP64)
//   trip  = NB blocks, then one warp vote.
//
// `cnt` folds the iteration budget into the escape filter: cnt = filter_hi + (B-1) - left, so cnt >= filter_hi exactly
// when fewer than B iterations are left (filter_hi >= hi(32.0) > 2^30 > iterations, so cnt never wraps).  A lane that
// stops keeps nothing but the position it had at the START of that block (the block's input registers are still live
// at the branch: no moves on the hot path) and its iteration index; the service code stores that pair as a
// provisional record into iters[]/zfinal[] and refills the lane.  K1r then re-runs at most B+1 iterations per pixel
// (2B-1 for a pixel that ran into the iteration cap) from the record with the reference's test after every step — at full lane efficiency, ~B/2 useful iterations per
// pixel against ~1400 in the frame — and overwrites the record with what calc::recursive returns.
//
// Why B+1 steps always suffice (same host-checked precondition as the sparse kernel, frb_api.cu: sparse_is_safe:
// R = limit >= 8, |c| <= R^2/2 - 3R, S = R^2 <= 1e300): a stop by the filter means the block's last position has
// max(re^2, im^2) >= S/2.  Either some step of the block escaped (an escaped orbit keeps growing to inf/NaN, which
// also fire the unsigned filter, so no escape is ever missed at a block end), or the last step was a near miss and
// the next one has |z| >= |z_prev|^2 - |c| >= 3R, which escapes.  A stop by the budget leaves r = iterations mod B < B
// iterations after the block just run, whose start is what the record holds: B + r <= 2B - 1 steps.
#pragma once

namespace frb {

#ifdef FRB_STREAM_TRACE
__device__ unsigned int g_warp_trips[16384];
__device__ unsigned int g_warp_smid[16384];
__device__ unsigned long long g_warp_exit[16384];
#endif

constexpr uint32_t kNoPixel = 0xffffffffu;

// B iterations of the 6-instruction form; nothing recorded.
template <int B>
__device__ __forceinline__ void deferred_block(double &x, double &y, double &a, double &b, const double cr, const double ci)
{
#pragma unroll
    for (int k = 0; k < B; ++k) {
        const double s = __dsub_rn(a, b);
        const double q = __dmul_rn(x, y);
        x = __dadd_rn(s, cr);
        y = __fma_rn(q, 2.0, ci);
        a = __dmul_rn(x, x);
        b = __dmul_rn(y, y);
    }
}

__device__ __forceinline__ uint32_t max3u(uint32_t a, uint32_t b, uint32_t c) { return max(max(a, b), c); }

// One trip of up to NB blocks.  Returns false for a lane that stopped.  The stopping lane writes its provisional record
// (iterations completed at the start of the stopping block, position there) from the cold side of the branch: the
// block's input registers are still live there, so the hot path carries no copies, and (a, b, cnt) keep the values that
// fired the filter so that the service code recognises the lane as stopped.
template <int B, int NB, int BLK = 0>
__device__ __forceinline__ bool deferred_trip(double &x, double &y, double &a, double &b, uint32_t &cnt, const double cr,
                                              const double ci, const uint32_t filter_hi, uint32_t *__restrict__ rec_iters,
                                              double2 *__restrict__ rec_z, const uint32_t rec_bias)
{
    if constexpr (BLK == NB) {
        return true;
    } else {
        const double x0 = x, y0 = y;
        deferred_block<B>(x, y, a, b, cr, ci);
        cnt += (uint32_t)B;
        if (__builtin_expect(max3u((uint32_t)__double2hiint(a), (uint32_t)__double2hiint(b), cnt) >= filter_hi, 0)) {
            // cnt = filter_hi + (B-1) - left_after and left_at_block_start = left_after + B.
            // The two stores sit in a (never repeated) inline-PTX loop, one copy per block: ptxas does not if-convert a
            // loop, so the stop path stays out of line and the hot path carries neither predicated stores nor the
            // register copies a merged stop path would need to bring every block's (x0, y0) into the same registers.
            asm volatile(
                "{\n\t.reg .pred q;\n"
                "FRB_STOP_%=:\n\t"
                "st.global.u32 [%0], %1;\n\t"
                "st.global.v2.f64 [%2], {%3, %4};\n\t"
                "setp.eq.u32 q, %1, 0xffffffff;\n\t"  // a record is an iteration index < 2^30
                "@q bra FRB_STOP_%=;\n\t"
                "}" ::"l"(rec_iters), "r"(cnt + rec_bias), "l"(rec_z), "d"(x0), "d"(y0), "n"(BLK)
                : "memory");
            return false;
        }
        return deferred_trip<B, NB, BLK + 1>(x, y, a, b, cnt, cr, ci, filter_hi, rec_iters, rec_z, rec_bias);
    }
}

// CT threads per CTA, MINB resident CTAs per SM.  Warps are independent workers, so the CTA shape only matters to the warp
// scheduler — and there it matters a lot: measured on B200 (scripts/trace_warps.py), warps of co-resident CTAs are NOT served
// evenly.  With 3 CTAs of 8 warps per SM the oldest CTA's warps made ~4000 trips per frame, the second's ~1250 and the
// youngest's ~150: two warps per SMSP already take 3/4 of the FP64 pipe and the last CTA starves (its pixels take
// milliseconds, which stretches the end of the frame and holds back the streaming finisher's watermarks).  Within one CTA
// the warps progress evenly, so the default shape is ONE wide CTA per SM.
template <bool JULIA, int B, int NB, int MINB, int CT>
__global__ void __launch_bounds__(CT, MINB)
frb_iterate_deferred_kernel(const __grid_constant__ RenderParams p)
{
    constexpr unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const uint32_t filter_hi = p.filter_hi;

    // Per-lane orbit.  A lane without a pixel (pix == kNoPixel) runs the zero orbit with cnt counting up from 0: it
    // cannot reach filter_hi (> 2^30) before the warp's other lanes exhaust their budgets (< 2^30).
    double x = 0.0, y = 0.0, a = 0.0, b = 0.0, cr = 0.0, ci = 0.0;
    uint32_t cnt = 0u;
    uint32_t pix = kNoPixel;

    // warp-uniform cursor into the current chunk (a segment of one image row) and the prefetched id of the next
    uint32_t cur_x = 0, cur_end = 0, cur_rowpix = 0;
    double cur_im = 0.0;
    bool drained = false;
    uint32_t next_chunk = 0;
    if (lane == 0) next_chunk = atomicAdd(p.queue, 1u);
    uint32_t trips = 0;  // warp-uniform
#ifdef FRB_STREAM_TRACE
    if (lane == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMin(p.stats + 13, t);
    }
#endif
    // Watermark for the streaming finisher (frb_stream.cuh): the lowest pixel index this warp still owes a record for.
    // Published (release) whenever the warp fetched a chunk or the queue is drained, and as 0xffffffff on exit.  A stale
    // value is always a lower bound: the queue hands out increasing chunk ids, so whatever the warp takes next lies above
    // everything it holds now.
    uint32_t *const wm = p.wm ? p.wm + (blockIdx.x * (CT / 32) + (threadIdx.x >> 5)) : nullptr;
    bool fetched = false;

    for (;;) {
        // ------------------------------ service (cold) ------------------------------
        // a stopped lane still holds the (a, b, cnt) that fired the filter; running lanes and idle lanes do not fire it
        if (max3u((uint32_t)__double2hiint(a), (uint32_t)__double2hiint(b), cnt) >= filter_hi) pix = kNoPixel;  // record already written
        unsigned need = __ballot_sync(FULL, pix == kNoPixel);
        if (need != 0u) {
            const uint32_t cnt0 = filter_hi + (uint32_t)(B - 1) - p.iterations;
            if (!drained) {
                const uint32_t n = (uint32_t)__popc(need);
                if (cur_end - cur_x >= n) {
                    // fast path: the warp's current row segment covers every free lane
                    if (pix == kNoPixel) {
                        const uint32_t px = cur_x + (uint32_t)__popc(need & lt_mask);
                        const double re0 = __ldg(p.tab_re + px);
                        // get_recursive_pixel (calc/src/lib.rs:208-210): Mandelbrot c = start, Julia c = julia_set
                        x = re0;
                        y = cur_im;
                        cr = JULIA ? p.julia_re : re0;
                        ci = JULIA ? p.julia_im : cur_im;
                        a = __dmul_rn(re0, re0);
                        b = __dmul_rn(cur_im, cur_im);
                        cnt = cnt0;
                        pix = cur_rowpix + px;
                    }
                    cur_x += n;
                    need = 0u;
                } else {
                    // generic path: spans a segment boundary and/or fetches the next chunk from the queue
                    while (need != 0u) {
                        if (cur_x == cur_end) {
                            const uint32_t c = __shfl_sync(FULL, next_chunk, 0);
                            if (c >= p.total_chunks) {
                                drained = true;
#ifdef FRB_STREAM_TRACE
                                if (lane == 0) {
                                    unsigned long long t;
                                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                                    atomicMin(p.stats + 9, t);
                                }
#endif
                                break;
                            }
                            if (lane == 0) next_chunk = atomicAdd(p.queue, 1u);
                            fetched = true;
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
                        if (pix == kNoPixel && rank < avail) {
                            const uint32_t px = cur_x + rank;
                            const double re0 = __ldg(p.tab_re + px);
                            x = re0;
                            y = cur_im;
                            cr = JULIA ? p.julia_re : re0;
                            ci = JULIA ? p.julia_im : cur_im;
                            a = __dmul_rn(re0, re0);
                            b = __dmul_rn(cur_im, cur_im);
                            cnt = cnt0;
                            pix = cur_rowpix + px;
                        }
                        cur_x += min((uint32_t)__popc(need), avail);
                        need = __ballot_sync(FULL, pix == kNoPixel);
                    }
                }
            }
            if (need != 0u) {
                // queue drained: lanes without a pixel park on the zero orbit; the warp leaves when all are parked
                if (need == FULL) break;
                if (pix == kNoPixel) {
                    x = y = a = b = cr = ci = 0.0;
                    cnt = 0u;
                }
            }
        }

        if (wm != nullptr && (fetched || drained)) {
            uint32_t low = __reduce_min_sync(FULL, pix);  // lanes without a pixel hold kNoPixel = 0xffffffff
            if (cur_x < cur_end) low = min(low, cur_rowpix + cur_x);
            __threadfence();  // every lane's records before the watermark
            __syncwarp();
            if (lane == 0) *reinterpret_cast<volatile uint32_t *>(wm) = low;
            fetched = false;
        }

        // ------------------------------ hot ------------------------------
        // Every lane enters with at least B iterations left (host: iterations >= B).  Trips repeat until some lane stopped.
        uint32_t *const rec_iters = p.iters + pix;  // parked lanes never stop, so their (wild) addresses are never used
        double2 *const rec_z = p.zfinal + pix;
        const uint32_t rec_bias = p.iterations - filter_hi - (uint32_t)(2 * B - 1);
        bool live;
        do {
            ++trips;
            live = deferred_trip<B, NB>(x, y, a, b, cnt, cr, ci, filter_hi, rec_iters, rec_z, rec_bias);
        } while (__all_sync(FULL, live));
    }

    if (wm != nullptr) {
        __threadfence();
        __syncwarp();
        if (lane == 0) *reinterpret_cast<volatile uint32_t *>(wm) = 0xffffffffu;
    }
    {
#ifdef FRB_STREAM_TRACE
        if (lane == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            atomicMax(p.stats + 6, t);
            const unsigned w = blockIdx.x * (CT / 32) + (threadIdx.x >> 5);
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            g_warp_trips[w] = trips;
            g_warp_smid[w] = smid;
            g_warp_exit[w] = t;
        }
#endif
    }
    // lane-slot iterations issued: every trip is counted at its full length (a trip all of whose lanes stop early is shorter)
    if (lane == 0) atomicAdd(p.stats + 1, (unsigned long long)trips * (unsigned long long)(32 * B * NB));
}

// ---------------------------------------------------------------------------------------------------------
// K1r: resolve.  One thread per pixel of the rows [p.resolve_row0, p.resolve_row0 + p.resolve_rows) of the shard:
// reads the provisional record (iterations completed, position), runs at most 2 * p.defer_block - 1 iterations with the
// reference's escape test after every step (calc/src/lib.rs:249-254, on the carried squares: same products, same
// sum, same compare) and overwrites the record with calc::recursive's return value: (next, i) on escape,
// (previous, iterations) otherwise.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void resolve_record(const RenderParams &p, uint32_t &count, double2 &z, const double cr, const double ci)
{
    const uint32_t iterations = p.iterations;
    const uint32_t done = count;
    // A stop by the filter resolves within B + 1 steps (header).  A stop by the budget happens after the block that leaves
    // r = iterations mod B < B iterations, and the record is that block's START: B + r <= 2B - 1 steps to the cap.
    const uint32_t steps = min(2u * p.defer_block - 1u, iterations - done);
    Orbit o = {z.x, z.y, __dmul_rn(z.x, z.x), __dmul_rn(z.y, z.y), cr, ci};
    const double S = p.limit_sq;
    const uint32_t filter_hi = p.filter_hi;
    count = iterations;
    // (Splitting this into a filter-only inner loop with the FP64 compare outside it was measured slower: C2 0.205 vs 0.177 ms.)
    for (uint32_t k = 0; k < steps; ++k) {
        stream_core<6>(o);
        if (stream_filter(o, filter_hi) && stream_escaped(o, S)) {
            count = done + k;
            break;
        }
    }
    z = make_double2(o.x, o.y);
}

__global__ void __launch_bounds__(256)
frb_resolve_kernel(const __grid_constant__ RenderParams p)
{
    const uint32_t x = blockIdx.y * blockDim.x + threadIdx.x;
    const uint32_t lr = p.resolve_row0 + blockIdx.x;
    if (x >= p.width) return;
    const uint32_t pix = lr * p.width + x;
    uint32_t count = p.iters[pix];
    double2 z = p.zfinal[pix];
    const bool julia = (p.algo == kAlgoJulia);
    const double cr = julia ? p.julia_re : __ldg(p.tab_re + x);
    const double ci = julia ? p.julia_im : __ldg(p.tab_im + local_row_to_y(p, lr));
    resolve_record(p, count, z, cr, ci);
    p.iters[pix] = count;
    p.zfinal[pix] = z;
}

}  // namespace frb
