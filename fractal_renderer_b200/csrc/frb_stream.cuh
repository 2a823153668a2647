// frb_stream.cuh — K2s: the streaming finisher.  Runs BESIDE the deferred iterate kernel (K1c) on a second stream, one CTA per
// SM in the CTA slot K1c leaves free, and shades the frame in row slabs as soon as each slab's pixels are recorded:
// provisional records -> resolve -> colour -> 16-byte stores straight into the destination, which for the host-output path
// (frb_render / frb_render_shard with a pinned buffer) is the caller's own memory, written over PCIe while K1c is still
// iterating the rest of the frame.  What stays exposed at the end of a frame is one small slab instead of a device->host
// copy of the whole image and of per-slab kernel launches (C2: 5.25 -> see profiles/r1_tuning_sweeps.txt).
//
// Completion of a slab is detected without any per-pixel accounting.  K1c's warps publish a watermark — the lowest pixel
// index they still owe a record for (frb_deferred.cuh) — after a __threadfence, whenever they fetch a chunk, on every
// service call once the queue is drained, and 0xffffffff on exit.  The queue hands out chunks in increasing order, so
//      min over all warps of watermark >= end of slab   =>   every pixel of the slab has its record in memory:
// a warp whose watermark is past the slab has taken a chunk beyond it, hence every chunk of the slab was handed out, and each
// of those is held by a warp whose watermark cannot pass its pixels before they are recorded.  Stale watermarks are lower
// bounds (a warp's next chunk always lies above everything it holds), slots start at 0, and warps that start late (grid
// larger than the free CTA slots) simply hold slabs back until they run — no co-residency is required, so nothing can
// deadlock; the finisher's own CTAs take their blocks from an atomic counter for the same reason (which of them are
// resident early is the block scheduler's choice).  As a last resort the wait is bounded and reports failure through ctl.status instead of hanging the GPU.
#pragma once

namespace frb {

struct StreamCtl {
    const uint32_t *wm;   // [nwarps] watermarks written by K1c
    uint32_t *words;      // [0] status: 0 = ok, 1 = a wait ran out (read by the host)   [1] min over wm, refreshed by the watcher
    uint32_t *queue;      // [0] watcher election (0 = vacant, else CTA index + 1)        [1] next work item
    const uint32_t *gen;  // generation word: K0 stores gen_value there once the words above and the watermarks are reset
    uint32_t *started;    // [0] number of finisher CTAs that saw the frame start (reset by K0); 0 = the host must shade itself
    uint32_t gen_value;
    uint32_t gen_polls;   // bound on the wait for the generation word (polls of ~1 us)
    uint32_t nwarps;
    uint32_t slab_rows;   // multiple of 16 (keeps every slab's first output byte 16-byte aligned)
    uint32_t max_polls;   // bound on one wait (polls of >= 2 us)
    uint32_t pad;
    uint32_t *slab_count; // nullptr, or [nslabs] device counters (zeroed by the host): work items of the slab already shaded
    uint32_t *slab_ready; // [nslabs] flags in mapped HOST memory: set by the CTA that shaded the slab's last item, after a system
                          // fence — the slab's pixels are in the (host) destination; host copy threads wait on them
    uint32_t blocks_per_slab;
    uint32_t pad2;
};

__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

#ifdef FRB_STREAM_TRACE
__device__ unsigned long long g_trace[3 * 8192];
__device__ unsigned int g_trace_n;
#endif

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t *p) { return *reinterpret_cast<const volatile uint32_t *>(p); }

// Blocks until min(watermarks) >= target.  The CTA that won the election (`watcher`) scans the watermark table and publishes
// the minimum; the others poll that one word and scan for themselves only every 32nd poll (the watcher may be busy shading).
// Which CTAs of this kernel are resident early is up to the block scheduler, hence an election rather than "CTA 0".
// Returns false (CTA-uniformly) when the wait ran out or another CTA reported failure.  s[0] = scratch minimum (kept at
// 0xffffffff between calls), s[1] = broadcast of the decision.
__device__ __forceinline__ bool stream_wait(const StreamCtl &ctl, uint32_t target, uint32_t *s, bool watcher)
{
    for (uint32_t poll = 0;; ++poll) {
        const bool scan = watcher || (poll & 31u) == 31u;
        if (scan) {
            uint32_t m = 0xffffffffu;
            for (uint32_t i = threadIdx.x; i < ctl.nwarps; i += blockDim.x) m = min(m, ld_volatile_u32(ctl.wm + i));
            m = __reduce_min_sync(0xffffffffu, m);
            if ((threadIdx.x & 31u) == 0) atomicMin(s, m);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t m;
            if (scan) {
                m = s[0];
                s[0] = 0xffffffffu;
                atomicMax(ctl.words + 1, m);  // watermarks only rise, so the freshest scan is the largest
#ifdef FRB_STREAM_TRACE
                if (watcher) {
                    const unsigned k = atomicAdd(&g_trace_n, 1u);
                    if (k < 8192u) {
                        const uint32_t arg = 0;
                        g_trace[3 * k] = globaltimer_ns();
                        g_trace[3 * k + 1] = m;
                        g_trace[3 * k + 2] = arg;
                    }
                }
#endif
            } else {
                m = ld_volatile_u32(ctl.words + 1);
            }
            const bool failed = ld_volatile_u32(ctl.words) != 0u;
            uint32_t verdict = 0u;  // 0 = keep waiting, 1 = go, 2 = give up
            if (failed || poll >= ctl.max_polls) verdict = 2u;
            else if (m >= target) verdict = 1u;
            if (verdict == 2u) *reinterpret_cast<volatile uint32_t *>(ctl.words) = 1u;
            s[1] = verdict;
        }
        __syncthreads();
        const uint32_t verdict = s[1];
        if (verdict == 1u) {
            __threadfence();  // acquire: the records behind the watermarks
            return true;
        }
        if (verdict == 2u) return false;
        __nanosleep(watcher ? 500u : 2000u);
    }
}

// Work items are the 1024-pixel shade blocks of the frame in slab order, handed out by one atomic counter: a CTA that only
// becomes resident late (or never, before K1c ends) does not sit on a share of the early slabs.
__global__ void __launch_bounds__(kShadeThreads)
frb_finish_stream_kernel(const __grid_constant__ RenderParams p, const __grid_constant__ StreamCtl ctl)
{
    __shared__ __align__(16) uint32_t stage[kShadeThreads * 3];
    __shared__ uint32_t s_wait[2];
    __shared__ uint32_t s_item[2];
    __shared__ double2 lg[128];
    load_log2_table(lg);
    if (threadIdx.x == 0) {
        // This kernel is launched BEFORE the frame's table kernel and K1c, with no dependency on them: the warp scheduler
        // favours the warps of older CTAs (frb_deferred.cuh), and the finisher — asleep most of the time, 3-4 % of the frame's
        // work — must not starve behind K1c's saturating FP64 stream.  It sleeps here until K0 has reset the control words.
        uint32_t ok = 0u;
        for (uint32_t poll = 0; poll < ctl.gen_polls; ++poll) {
            if (ld_volatile_u32(ctl.gen) == ctl.gen_value) {
                ok = 1u;
                break;
            }
            __nanosleep(1000u);
        }
        // A CTA that never sees the frame start (kernels serialised by a profiler, or the frame held up behind other work for
        // longer than the bound) just leaves: blocks are handed out dynamically, so the CTAs that do start shade everything,
        // and if none does the host falls back to the ordinary shade kernel (frb_api.cu: render_blocking).
        if (ok) atomicAdd(ctl.started, 1u);
        __threadfence();
        s_wait[0] = 0xffffffffu;
        s_item[0] = ok;
        s_item[1] = (ok && atomicCAS(ctl.queue, 0u, blockIdx.x + 1u) == 0u) ? 1u : 0u;
    }
    __syncthreads();
    const bool started = s_item[0] != 0u;
    const bool watcher = s_item[1] != 0u;
    __syncthreads();
    if (!started) return;
#ifdef FRB_STREAM_TRACE
    if (threadIdx.x == 0) atomicMin(p.stats + 3, globaltimer_ns());
#endif
    const uint64_t cta_px = (uint64_t)kShadeThreads * kShadePxPerThread;
    const uint32_t nslabs = (p.rows + ctl.slab_rows - 1) / ctl.slab_rows;
    const uint32_t bps = (uint32_t)(((uint64_t)ctl.slab_rows * p.width + cta_px - 1) / cta_px);  // blocks per (full) slab
    const uint32_t nitems = nslabs * bps;
    unsigned long long useful = 0;
    uint32_t released = 0;  // slabs [0, released) are known complete
    for (;;) {
        if (threadIdx.x == 0) s_item[0] = atomicAdd(ctl.queue + 1, 1u);
        __syncthreads();
        const uint32_t item = s_item[0];
        __syncthreads();
        if (item >= nitems) break;
        const uint32_t slab = item / bps, blk = item - slab * bps;
        const uint32_t r0 = slab * ctl.slab_rows, r1 = min(p.rows, r0 + ctl.slab_rows);
        if (slab >= released) {
            // pixel indices are < 2^32 - 16 (frb_api.cu: make_plan), so the end of the last slab still compares below 0xffffffff
            if (!stream_wait(ctl, r1 * p.width, s_wait, watcher)) break;
            released = slab + 1;
#ifdef FRB_STREAM_TRACE
            if (threadIdx.x == 0 && slab == 0) atomicMin(p.stats + 4, globaltimer_ns());
            if (threadIdx.x == 0 && slab == nslabs - 1) atomicMin(p.stats + 7, globaltimer_ns());
#endif
        }
        RenderParams sh = p;  // the slab as a little image of its own (the shade code indexes from its pointers)
        const size_t pix0 = (size_t)r0 * p.width;
        sh.rows = r1 - r0;
        sh.resolve_row0 = p.resolve_row0 + r0;
        sh.iters = p.iters + pix0;
        sh.zfinal = p.zfinal + pix0;
        offset_out(sh, p, pix0);
        shade_blocks(sh, stage, lg, blk, 0xffffffffffull, useful);  // exactly one block (none if blk is past a short last slab)
        if (ctl.slab_count != nullptr) {
            // destination = the library's pinned staging buffer (frb_api.cu: pageable destinations): tell the host's copy threads
            // when the slab is complete.  Counting happens in device memory (atomics on mapped host memory are not something to
            // lean on); only the plain "ready" store crosses PCIe, behind the pixels of every CTA that counted before this one.
            __threadfence_system();  // this thread's pixels before the count
            __syncthreads();
            if (threadIdx.x == 0 && atomicAdd(ctl.slab_count + slab, 1u) + 1u == ctl.blocks_per_slab) {
                __threadfence_system();
                *reinterpret_cast<volatile uint32_t *>(ctl.slab_ready + slab) = 1u;
            }
        }
    }
    shade_publish_useful(p, useful);
#ifdef FRB_STREAM_TRACE
    if (threadIdx.x == 0) atomicMax(p.stats + 5, globaltimer_ns());
#endif
}

}  // namespace frb
