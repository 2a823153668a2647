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
// Why B+1 steps always suffice (host-checked precondition, frb_api.cu: deferred_is_safe:
// 8 <= R = limit <= 2^20, |c| <= R^2/2 - 3R): a stop by the filter means the block's last position has
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

// ---- the end of a frame ------------------------------------------------------------------------------------------------
// Measured (profiles/r2_probe1_timeline_latency_pageable.txt): from the moment the queue runs dry to the exit of the last warp
// a C2 frame spends 0.23 ms — whatever the size of the frame or shard, i.e. 5 % of the whole frame and 30 % of an eighth of
// it — with the FP64 pipe draining all the while.  Two causes, two remedies:
//  * a warp holds up to a chunk of unstarted pixels plus a prefetched chunk when the queue runs dry, each possibly a
//    full-length orbit.  -> The last rows of the launch are handed out in SHORT chunks (RenderParams::q_chunk2), so that what
//    a warp still holds at that moment is a few pixels, not 64.
//  * after that, every warp keeps running until its slowest lane is done, more and more of its lanes parked: 24 warps per SM
//    at a few live lanes each share the pipe at 72 cycles per iteration, where 2 full warps per SM sub-partition would run the
//    same orbits at the 24-cycle latency bound of the dependent chain (scripts/latency_probe.cu).  -> Every SM sub-partition
//    has an ORBIT POOL in shared memory: a drained warp left with at most q_pool_max live lanes pushes their orbits (position,
//    carried squares, c, budget counter, pixel) into the pool of its sub-partition and exits; drained warps with parked lanes
//    pop them.  Live lanes thus collect in few, full warps, evenly over the four schedulers.  A warp only donates when the
//    parked lanes registered by the other warps of its sub-partition can take all its orbits at once — an orbit waiting in
//    the pool makes no progress at all (first version, one pool per CTA without that rule: the tail grew from 0.18 to 0.27 ms,
//    profiles/r2_pool_sweep1_cta_pool_rejected.txt).  All pool traffic is on the cold side, under a spin lock, after the queue is dry.
//    Measured with both (profiles/r2_pool_sweep2_smsp_pools.txt): queue dry -> last warp 0.23 -> 0.13 ms, median SM 0.16 -> 0.10 ms.
//    Also tried, no gain: retiring all but 2-3 warps per sub-partition for the last 25-400 Kpx of the frame so that late pixels
//    run at the chain's latency bound from the start (C2 1/8: 0.716 -> 0.713-0.728 ms).
constexpr int kPoolCap = 160;  // per sub-partition: >= (warps per sub-partition - 1) x the largest q_pool_max (31)

struct OrbitPool {
    uint32_t lock;               // 0 = free
    uint32_t count;              // orbits in the pool (a stack)
    uint32_t alive;              // warps of this sub-partition that have neither exited nor donated their lanes
    uint32_t free;               // parked lanes registered by the drained warps of this sub-partition
    double x[kPoolCap], y[kPoolCap], a[kPoolCap], b[kPoolCap], cr[kPoolCap], ci[kPoolCap];
    uint32_t cnt[kPoolCap], pix[kPoolCap], donor[kPoolCap];
};

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
    __shared__ OrbitPool pools[4];
    __shared__ uint32_t pool_remaining[32];  // per warp: pixels it took from the queue itself and donated, not yet recorded
    const unsigned lane = threadIdx.x & 31u;
    const unsigned warp_in_cta = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const uint32_t filter_hi = p.filter_hi;
    const uint32_t pool_max = p.q_pool_max;  // 0: no pool
    // warps w, w + 4, w + 8, ... of a CTA run on the same SM sub-partition (scheduler): they share a pool
    OrbitPool &pool = pools[warp_in_cta & 3u];
    if (pool_max != 0u) {
        if (threadIdx.x < 32u) pool_remaining[threadIdx.x] = 0u;
        if (threadIdx.x < 4u) {
            pools[threadIdx.x].lock = 0u;
            pools[threadIdx.x].count = 0u;
            pools[threadIdx.x].free = 0u;
            pools[threadIdx.x].alive = (CT / 32 + 3 - threadIdx.x) / 4;
        }
        __syncthreads();  // the only CTA-wide barrier of the kernel: before any warp has work
    }
    uint32_t my_free = 0u;  // warp-uniform: the parked lanes this warp has registered in pool.free

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
    // everything it holds now.  Pixels never change watermark: an orbit donated to the pool stays covered by the warp that
    // took it from the queue (its `origin`), which does not publish its exit before every such pixel has its record — the
    // finisher's scan of the table is not atomic, so coverage must never move between slots.
    uint32_t *const wm = p.wm ? p.wm + (blockIdx.x * (CT / 32) + warp_in_cta) : nullptr;
    bool fetched = false;
    uint32_t origin = kNoPixel;  // per lane: the warp (in this CTA) whose pixel this lane runs after taking it from the pool
    bool poll_pool = false;  // warp-uniform: drained, lanes parked, so look into the pool after every trip

    for (;;) {
        // ------------------------------ service (cold) ------------------------------
        // a stopped lane still holds the (a, b, cnt) that fired the filter; running lanes and idle lanes do not fire it
        if (max3u((uint32_t)__double2hiint(a), (uint32_t)__double2hiint(b), cnt) >= filter_hi) {
            pix = kNoPixel;  // record already written
            if (origin != kNoPixel) {
                // a pixel taken over from the pool: tell the warp that owes it to the finisher, after the record
                if (wm != nullptr) {
                    __threadfence();
                    atomicSub(&pool_remaining[origin], 1u);
                }
                origin = kNoPixel;
            }
        }
        unsigned need = __ballot_sync(FULL, pix == kNoPixel);
        bool leave = false;
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
                            // chunk -> row segment: long chunks first, the short chunks of the end phase after them
                            uint32_t lr, x0, len;
                            if (c < p.q_nbig) {
                                const uint32_t r = c / p.chunks_per_row;
                                lr = p.q_row0 + r;
                                x0 = (c - r * p.chunks_per_row) * p.chunk;
                                len = p.chunk;
                            } else {
                                const uint32_t c2 = c - p.q_nbig;
                                const uint32_t r = c2 / p.q_cpr2;
                                lr = p.q_row0 + p.q_rows_big + r;
                                x0 = (c2 - r * p.q_cpr2) * p.q_chunk2;
                                len = p.q_chunk2;
                            }
                            cur_x = x0;
                            cur_end = min(x0 + len, p.width);
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
                // Queue drained.
                if (pool_max == 0u) {
                    if (need == FULL) break;  // no lane holds a pixel: the warp leaves
                } else {
                    // One decision per service call, taken by lane 0 under the pool's lock: take orbits over, leave, or donate.
                    const uint32_t nfree = (uint32_t)__popc(need), live = 32u - nfree;
                    uint32_t action = 0u, k = 0u, base = 0u;  // action: 0 keep going, 1 take k orbits at base, 2 exit, 3 donate at base
                    uint32_t reg = nfree;                      // parked lanes this warp has registered after the decision
                    if (lane == 0) {
                        while (atomicCAS(&pool.lock, 0u, 1u) != 0u) __nanosleep(32);
                        const uint32_t have = *reinterpret_cast<volatile uint32_t *>(&pool.count);
                        const uint32_t alive = *reinterpret_cast<volatile uint32_t *>(&pool.alive);
                        const uint32_t fr = *reinterpret_cast<volatile uint32_t *>(&pool.free);
                        if (have != 0u) {
                            action = 1u;
                            k = min(have, nfree);
                            base = have - k;
                            reg = nfree - k;
                            *reinterpret_cast<volatile uint32_t *>(&pool.count) = base;
                        } else if (live == 0u) {
                            action = 2u;
                            reg = 0u;
                            *reinterpret_cast<volatile uint32_t *>(&pool.alive) = alive - 1u;
                        } else if (live <= pool_max && alive > 1u && live <= fr - my_free && live <= (uint32_t)kPoolCap) {
                            // the parked lanes of the OTHER warps of this sub-partition take all of this warp's orbits at once
                            action = 3u;
                            base = 0u;
                            reg = 0u;
                            *reinterpret_cast<volatile uint32_t *>(&pool.count) = live;
                            *reinterpret_cast<volatile uint32_t *>(&pool.alive) = alive - 1u;
                        }
                        *reinterpret_cast<volatile uint32_t *>(&pool.free) = fr - my_free + reg;
                    }
                    my_free = __shfl_sync(FULL, reg, 0);
                    action = __shfl_sync(FULL, action, 0);
                    k = __shfl_sync(FULL, k, 0);
                    base = __shfl_sync(FULL, base, 0);
                    if (action == 1u) {
                        const uint32_t rank = (uint32_t)__popc(need & lt_mask);
                        if (pix == kNoPixel && rank < k) {
                            const uint32_t i = base + rank;
                            x = *reinterpret_cast<volatile double *>(&pool.x[i]);
                            y = *reinterpret_cast<volatile double *>(&pool.y[i]);
                            a = *reinterpret_cast<volatile double *>(&pool.a[i]);
                            b = *reinterpret_cast<volatile double *>(&pool.b[i]);
                            cr = *reinterpret_cast<volatile double *>(&pool.cr[i]);
                            ci = *reinterpret_cast<volatile double *>(&pool.ci[i]);
                            cnt = *reinterpret_cast<volatile uint32_t *>(&pool.cnt[i]);
                            pix = *reinterpret_cast<volatile uint32_t *>(&pool.pix[i]);
                            origin = *reinterpret_cast<volatile uint32_t *>(&pool.donor[i]);
                        }
                    } else if (action == 3u) {
                        if (pix != kNoPixel) {
                            const uint32_t i = base + (uint32_t)__popc(~need & lt_mask);
                            *reinterpret_cast<volatile double *>(&pool.x[i]) = x;
                            *reinterpret_cast<volatile double *>(&pool.y[i]) = y;
                            *reinterpret_cast<volatile double *>(&pool.a[i]) = a;
                            *reinterpret_cast<volatile double *>(&pool.b[i]) = b;
                            *reinterpret_cast<volatile double *>(&pool.cr[i]) = cr;
                            *reinterpret_cast<volatile double *>(&pool.ci[i]) = ci;
                            *reinterpret_cast<volatile uint32_t *>(&pool.cnt[i]) = cnt;
                            *reinterpret_cast<volatile uint32_t *>(&pool.pix[i]) = pix;
                            *reinterpret_cast<volatile uint32_t *>(&pool.donor[i]) = origin != kNoPixel ? origin : warp_in_cta;
                        }
                        if (wm != nullptr) {
                            // pixels this warp took from the queue itself stay its responsibility until they are recorded
                            const uint32_t own = (uint32_t)__popc(__ballot_sync(FULL, pix != kNoPixel && origin == kNoPixel));
                            if (lane == 0 && own != 0u) atomicAdd(&pool_remaining[warp_in_cta], own);
                        }
                    }
                    __threadfence_block();
                    __syncwarp();
                    if (lane == 0) atomicExch(&pool.lock, 0u);
                    leave = action == 2u || action == 3u;
                    need = __ballot_sync(FULL, pix == kNoPixel);
                    if (action == 3u) need = FULL;
                }
                // lanes without a pixel park on the zero orbit
                if (!leave && pix == kNoPixel) {
                    x = y = a = b = cr = ci = 0.0;
                    cnt = 0u;
                }
            }
        }
        if (leave) {
            if (wm != nullptr) {
                // pixels this warp donated (now or earlier, as a taker that donated on) stay under its last watermark until
                // whoever runs them has written their records
                while (*reinterpret_cast<volatile uint32_t *>(&pool_remaining[warp_in_cta]) != 0u) __nanosleep(200);
            }
            break;
        }
        poll_pool = pool_max != 0u && drained && need != 0u;

        if (wm != nullptr && (fetched || drained)) {
            uint32_t low = __reduce_min_sync(FULL, pix);  // lanes without a pixel hold kNoPixel = 0xffffffff
            if (cur_x < cur_end) low = min(low, cur_rowpix + cur_x);
            __threadfence();  // every lane's records before the watermark
            __syncwarp();
            if (lane == 0) *reinterpret_cast<volatile uint32_t *>(wm) = low;
            fetched = false;
        }

        // ------------------------------ hot ------------------------------
        // Every lane enters with at least B iterations left (host: iterations >= B).  Trips repeat until some lane stopped
        // (or, for a drained warp with parked lanes, until the pool has something to take over).
        uint32_t *const rec_iters = p.iters + pix;  // parked lanes never stop, so their (wild) addresses are never used
        double2 *const rec_z = p.zfinal + pix;
        const uint32_t rec_bias = p.iterations - filter_hi - (uint32_t)(2 * B - 1);
        bool live;
        if (!poll_pool) {
            do {
                ++trips;
                live = deferred_trip<B, NB>(x, y, a, b, cnt, cr, ci, filter_hi, rec_iters, rec_z, rec_bias);
            } while (__all_sync(FULL, live));
        } else {
            do {
                ++trips;
                live = deferred_trip<B, NB>(x, y, a, b, cnt, cr, ci, filter_hi, rec_iters, rec_z, rec_bias);
                if (*reinterpret_cast<volatile uint32_t *>(&pool.count) != 0u) live = false;  // same word for every lane: uniform
            } while (__all_sync(FULL, live));
        }
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
