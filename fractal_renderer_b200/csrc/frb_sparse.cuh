// frb_sparse.cuh — K1b, the default iterate kernel: persistent lane-refill warps with SPARSE escape checks and a
// multi-block trip.  Included by frb_kernels.cuh.
//
// Measured on B200 (scripts/issue_probe.cu, profiles/): an FP64 instruction occupies an SMSP's issue port for two
// cycles and every other instruction for one more, so every non-FP64 instruction in the loop is FP64 throughput
// lost.  This kernel therefore issues, per block of 8 iterations, the 48 FP64 instructions and 5 others:
//
//   trip  = up to NB blocks;
//   block = 8 x { DADD s=a-b ; DMUL q=x*y ; DADD x'=s+c.re ; DFMA y'=fma(q,2,c.im) ; DMUL a=x'^2 ; DMUL b=y'^2 }
//           IADD left-=8 ; VIMNMX m=max(hi(a),hi(b)) ; ISETP m>=filter ; ISETP.OR left<8 ; @P BRA trip_end
//   each step's position is written to its own registers hx[k], hy[k] (the block rotates through them, no moves);
//   a lane that branches out keeps the recorded positions of its last block and waits at the end of the trip.
//   After the trip one VOTE decides whether any lane needs the (cold) service code.
//   (Per-lane predication of the FP64 instructions instead of the branch was tried in inline PTX: ptxas turns every
//   predicated FP64 write into an unpredicated op plus two FSELs, which is far worse.)
//
// Why one test per block suffices (host-checked precondition, frb_api.cu: sparse_is_safe): with R = limit >= 8
// and |c| <= R^2/2 - 3R for every pixel, a step with max(re^2, im^2) >= S/2 (S = R^2) is followed by
// |z_m| >= |z_{m-1}|^2 - |c| >= 3R, i.e. by escaped steps only, growing doubly exponentially to +inf and then NaN
// (inf - inf) — all of which fire the unsigned high-word filter.  So "filter silent at the end of the block"
// proves no step of the block escaped, and "first step that passes the necessary condition" is monotone in k.
// When the filter fires, that first step is found by an integer binary search over the 8 recorded positions and
// the exact reference test (re^2 + im^2 > S: same products, same sum, same compare) is evaluated on it (and on the
// next step if it was a near miss); the count and final position are bit-identical to calc::recursive.
#pragma once

namespace frb {

constexpr int kSparseU = 8;
// A lane is idle (holds no pixel) iff left >= kIdleBit; the host only selects this kernel for iterations < 2^30.
constexpr uint32_t kIdleBit = 0x40000000u;

__device__ __forceinline__ uint32_t abs_hi(double v) { return (uint32_t)__double2hiint(v) & 0x7fffffffu; }
__device__ __forceinline__ double sel_d(bool c, double a, double b) { return c ? a : b; }

// One block: 8 iterations of the 6-instruction form, each step's position left in its own registers.
// In: position (hx[7], hy[7]) and its squares (a, b).  Out: hx[0..7], hy[0..7] = positions after steps 0..7, (a, b) =
// squares of (hx[7], hy[7]).
__device__ __forceinline__ void sparse_block(double (&hx)[8], double (&hy)[8], double &a, double &b, const double cr, const double ci)
{
    double x = hx[7], y = hy[7];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const double s = __dsub_rn(a, b);
        const double q = __dmul_rn(x, y);
        x = __dadd_rn(s, cr);
        y = __fma_rn(q, 2.0, ci);
        a = __dmul_rn(x, x);
        b = __dmul_rn(y, y);
        hx[k] = x;
        hy[k] = y;
    }
}

// One trip of up to NB blocks.  A lane leaves the trip (keeping the recorded positions of its last block) at the
// first block whose end-of-block filter fires or that leaves fewer than 8 iterations; lanes that leave wait at the
// end of the trip.  Per block: 48 FP64 + IADD, VIMNMX, 2 x ISETP, BRA.  Returns true iff the lane ran all NB blocks and
// can run another.
template <int NB>
__device__ __forceinline__ bool sparse_trip(double (&hx)[8], double (&hy)[8], double &a, double &b, uint32_t &left,
                                            const double cr, const double ci, const uint32_t filter_hi)
{
#pragma unroll
    for (int blk = 0; blk < NB; ++blk) {
        sparse_block(hx, hy, a, b, cr, ci);
        left -= 8u;
        const uint32_t m = max((uint32_t)__double2hiint(a), (uint32_t)__double2hiint(b));
        if (m >= filter_hi || left < 8u) return false;
    }
    return true;
}

template <bool JULIA, int NB>
__global__ void __launch_bounds__(kIterateThreads, 4)
frb_iterate_sparse_kernel(const __grid_constant__ RenderParams p)
{
    constexpr unsigned FULL = 0xffffffffu;
    constexpr uint32_t U = (uint32_t)kSparseU;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;

    const double S = p.limit_sq;
    const uint32_t filter_hi = p.filter_hi;

    // Per-lane orbit: position (hx[7], hy[7]), its squares (a, b), constant (cr, ci).  A lane without a pixel runs
    // the zero orbit (never fires the filter) with left >= kIdleBit.  Every lane starts idle, so the first pass
    // through the service code is a pure refill.
    double hx[kSparseU], hy[kSparseU];
#pragma unroll
    for (int k = 0; k < kSparseU; ++k) hx[k] = hy[k] = 0.0;
    double a = 0.0, b = 0.0, cr = 0.0, ci = 0.0;
    uint32_t left = kIdleBit * 2u;
    uint32_t pix = 0;

    // warp-uniform cursor into the current chunk (a segment of one image row) and the prefetched id of the next
    uint32_t cur_x = 0, cur_end = 0, cur_rowpix = 0;
    double cur_im = 0.0;
    bool drained = false;
    uint32_t next_chunk = 0;
    if (lane == 0) next_chunk = atomicAdd(p.queue, 1u);

    unsigned long long lane_slots = 0;  // iterations issued per lane slot: the longest-running lane of every trip
    uint32_t left_mark = left;

    for (;;) {
        // ------------------------------ service (cold) ------------------------------
        lane_slots += (unsigned long long)__reduce_max_sync(FULL, left_mark - left);
        const uint32_t iterations = p.iterations;
        const bool cand = max((uint32_t)__double2hiint(a), (uint32_t)__double2hiint(b)) >= filter_hi;
        bool done = false;
        uint32_t count = iterations;
        double zx = hx[7], zy = hy[7];
        if (cand) {
            // First step k of the lane's last block whose position passes the necessary condition
            // max(|re|,|im|) >= sqrt(S/2): integer binary search, positions selected along the way.
            const uint32_t t2 = p.filter_abs_hi;
            const bool b2 = max(abs_hi(hx[3]), abs_hi(hy[3])) < t2;  // first pass is in 4..7
            const double x1 = sel_d(b2, hx[5], hx[1]), y1 = sel_d(b2, hy[5], hy[1]);
            const double x0 = sel_d(b2, hx[4], hx[0]), x2 = sel_d(b2, hx[6], hx[2]), x3 = sel_d(b2, hx[7], hx[3]);
            const double y0 = sel_d(b2, hy[4], hy[0]), y2 = sel_d(b2, hy[6], hy[2]), y3 = sel_d(b2, hy[7], hy[3]);
            const bool b1 = max(abs_hi(x1), abs_hi(y1)) < t2;        // ... in the upper pair of that half
            const double u0 = sel_d(b1, x2, x0), u1 = sel_d(b1, x3, x1);
            const double v0 = sel_d(b1, y2, y0), v1 = sel_d(b1, y3, y1);
            const bool b0 = max(abs_hi(u0), abs_hi(v0)) < t2;
            double ex = sel_d(b0, u1, u0), ey = sel_d(b0, v1, v0);
            uint32_t k = (b2 ? 4u : 0u) + (b1 ? 2u : 0u) + (b0 ? 1u : 0u);
            // the reference's test on that step (calc/src/lib.rs:250-251)
            bool esc = __dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey)) > S;
            if (!esc && k < U - 1) {
                // near miss (S/2 <= max < S without escaping): the next step is the first that can have escaped
                ++k;
                ex = hx[1]; ey = hy[1];
                if (k == 2) { ex = hx[2]; ey = hy[2]; }
                if (k == 3) { ex = hx[3]; ey = hy[3]; }
                if (k == 4) { ex = hx[4]; ey = hy[4]; }
                if (k == 5) { ex = hx[5]; ey = hy[5]; }
                if (k == 6) { ex = hx[6]; ey = hy[6]; }
                if (k == 7) { ex = hx[7]; ey = hy[7]; }
                esc = __dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey)) > S;
            }
            if (esc) {
                done = true;
                count = iterations - (left + U) + k;  // index of the block's first iteration + k
                zx = ex;
                zy = ey;
            }
        }
        if (!done && left < U) {
            // fewer than U iterations left (idle lanes have left >= kIdleBit): finish with the per-iteration test
            done = true;
            Orbit o = {hx[7], hy[7], a, b, cr, ci};
            const uint32_t i0 = iterations - left;
            for (uint32_t k = 0; k < left; ++k) {
                stream_core<6>(o);
                if (stream_filter(o, filter_hi) && stream_escaped(o, S)) {
                    count = i0 + k;
                    break;
                }
            }
            zx = o.x;
            zy = o.y;
        }
        if (done) {
            // recursive() returns (next, i) on escape, (previous, iterations) otherwise (calc/src/lib.rs:252-256)
            p.iters[pix] = count;
            p.zfinal[pix] = make_double2(zx, zy);
            left = kIdleBit * 2u;
        }
        unsigned need = __ballot_sync(FULL, left >= kIdleBit);
        if (need != 0u) {
            if (!drained) {
                const uint32_t n = (uint32_t)__popc(need);
                if (cur_end - cur_x >= n) {
                    // fast path: the warp's current row segment covers every free lane
                    if (left >= kIdleBit) {
                        const uint32_t px = cur_x + (uint32_t)__popc(need & lt_mask);
                        const double re0 = __ldg(p.tab_re + px);
                        // get_recursive_pixel (calc/src/lib.rs:208-210): Mandelbrot c = start, Julia c = julia_set
                        hx[7] = re0;
                        hy[7] = cur_im;
                        cr = JULIA ? p.julia_re : re0;
                        ci = JULIA ? p.julia_im : cur_im;
                        a = __dmul_rn(re0, re0);
                        b = __dmul_rn(cur_im, cur_im);
                        left = iterations;
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
                        if (left >= kIdleBit && rank < avail) {
                            const uint32_t px = cur_x + rank;
                            const double re0 = __ldg(p.tab_re + px);
                            hx[7] = re0;
                            hy[7] = cur_im;
                            cr = JULIA ? p.julia_re : re0;
                            ci = JULIA ? p.julia_im : cur_im;
                            a = __dmul_rn(re0, re0);
                            b = __dmul_rn(cur_im, cur_im);
                            left = iterations;
                            pix = cur_rowpix + px;
                        }
                        cur_x += min((uint32_t)__popc(need), avail);
                        need = __ballot_sync(FULL, left >= kIdleBit);
                    }
                }
            }
            if (need != 0u) {
                // queue drained: lanes without a pixel park on the zero orbit; the warp leaves when all are parked
                if (need == FULL) break;
                if (left >= kIdleBit) {
                    hx[7] = hy[7] = a = b = cr = ci = 0.0;
                    left = kIdleBit * 2u;
                }
            }
        }
        left_mark = left;

        // ------------------------------ hot ------------------------------
        // Every lane enters with left >= U (idle lanes: >= kIdleBit).  Trips repeat until some lane stopped.
        bool live;
        do {
            live = sparse_trip<NB>(hx, hy, a, b, left, cr, ci, filter_hi);
        } while (__all_sync(FULL, live));
    }

    if (lane == 0) atomicAdd(p.stats + 1, lane_slots * 32ull);
}

}  // namespace frb
