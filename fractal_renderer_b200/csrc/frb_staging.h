// frb_staging.h — pageable destinations: pinned staging buffers and the host copy threads.  Included by frb_api.cu inside its
// anonymous namespace (uses its Tuning, fail() and CUDA headers); not a stand-alone header.
#pragma once

// ---- pageable destinations: pinned staging + host copy threads ----------------------------------------------------------------
// The reference's get_image returns an owned Vec<RGB> (src/lib.rs:253): fresh, pageable, untouched memory that the GPU cannot
// write.  Copying into it AFTER the frame costs a staged cudaMemcpy plus one page fault per 4 KB — 2.1 ms on a 27 MB C2 frame,
// 11 ms on a 50 MB C3 frame (profiles/r2_probe1_timeline_latency_pageable.txt).  Instead the streaming finisher writes the image
// into a pinned staging buffer of the library, counts the finished work items of every slab in mapped host memory, and a few host
// threads copy each slab into the caller's buffer as soon as its count is full — first touch of the destination pages
// included, spread over the threads, all of it under the iterations of the rest of the frame.
constexpr size_t kMaxStageBytes = (size_t)3 << 30;  // larger shards take the sliced copy path

struct StageBuf {
    uint8_t *p = nullptr;       // pinned, portable, device-visible under UVA
    size_t cap = 0;             // image bytes
    uint32_t *done = nullptr;   // kMaxStageSlabs "slab ready" flags behind the image (written by the finisher)
    bool busy = false;
};

class StagePool {
    std::mutex mu_;
    std::vector<StageBuf *> bufs_;

public:
    StageBuf *acquire(size_t bytes)
    {
        std::lock_guard<std::mutex> lk(mu_);
        StageBuf *best = nullptr;
        for (StageBuf *b : bufs_)
            if (!b->busy && b->cap >= bytes && (!best || b->cap < best->cap)) best = b;
        if (!best) {
            // replace the SMALLEST idle buffer (every idle one is too small for this frame; the smallest is the least likely to
            // fit the next) rather than hoarding: at most 4 buffers are kept.  cudaFreeHost waits for the devices' work in
            // flight — this only happens when a process moves to a larger frame size for good.
            if (bufs_.size() >= 4) {
                size_t victim = bufs_.size();
                for (size_t i = 0; i < bufs_.size(); ++i)
                    if (!bufs_[i]->busy && (victim == bufs_.size() || bufs_[i]->cap < bufs_[victim]->cap)) victim = i;
                if (victim != bufs_.size()) {
                    cudaFreeHost(bufs_[victim]->p);
                    delete bufs_[victim];
                    bufs_.erase(bufs_.begin() + (long)victim);
                }
            }
            const size_t cap = bytes + bytes / 8 + (1u << 20);
            const size_t total = ((cap + 255) & ~(size_t)255) + kMaxStageSlabs * sizeof(uint32_t);
            void *q = nullptr;
            if (cudaHostAlloc(&q, total, cudaHostAllocPortable) != cudaSuccess) {
                cudaGetLastError();
                return nullptr;
            }
            best = new StageBuf();
            best->p = static_cast<uint8_t *>(q);
            best->cap = cap;
            best->done = reinterpret_cast<uint32_t *>(best->p + ((cap + 255) & ~(size_t)255));
            bufs_.push_back(best);
        }
        best->busy = true;
        return best;
    }
    void release(StageBuf *b)
    {
        std::lock_guard<std::mutex> lk(mu_);
        b->busy = false;
    }
    void clear()
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (StageBuf *b : bufs_) {
            cudaFreeHost(b->p);
            delete b;
        }
        bufs_.clear();
    }
};
StagePool g_stage;

// Asks the kernel to back a large destination with huge pages (THP in "madvise" mode, the default of most distributions): a fresh
// 50 MB Vec is 12 800 first-touch faults of 4 KB — or 25 of 2 MB.  A hint about memory the caller owns: it changes how the pages
// behind it are allocated, never their contents.  Failure (not anonymous memory, THP off) is ignored.
void advise_huge_pages(uint8_t *p, size_t bytes)
{
    if (!tuning().hugepage || bytes < ((size_t)4 << 20)) return;
    const uintptr_t page = 4096;
    const uintptr_t a = (reinterpret_cast<uintptr_t>(p) + page - 1) & ~(page - 1);
    const uintptr_t e = (reinterpret_cast<uintptr_t>(p) + bytes) & ~(page - 1);
    if (e > a) madvise(reinterpret_cast<void *>(a), e - a, MADV_HUGEPAGE);
}

// First touch of destination pages without changing what they hold.  MADV_POPULATE_WRITE (Linux 5.14) faults a whole range in
// one system call; older kernels get one atomic OR-with-zero per page, which cannot lose a byte another thread is copying into
// the same page.  Every page touched holds at least one byte of the caller's buffer.
#ifndef MADV_POPULATE_WRITE
#define MADV_POPULATE_WRITE 23
#endif
std::atomic<int> g_populate_by_madvise{1};

void populate_range(uint8_t *lo, uint8_t *hi)
{
    if (hi <= lo) return;
    const uintptr_t page = 4096;
    const uintptr_t a = reinterpret_cast<uintptr_t>(lo) & ~(page - 1);
    const uintptr_t e = (reinterpret_cast<uintptr_t>(hi) + page - 1) & ~(page - 1);
    if (g_populate_by_madvise.load(std::memory_order_relaxed)) {
        if (madvise(reinterpret_cast<void *>(a), e - a, MADV_POPULATE_WRITE) == 0) return;
        if (errno == EINVAL) g_populate_by_madvise.store(0, std::memory_order_relaxed);  // kernel without it
        else if (errno != EAGAIN && errno != EINTR) return;  // not ours to populate (file mapping without room, ...): the copy will fault it
    }
    for (uintptr_t q = a; q < e; q += page) {
        uint8_t *b = reinterpret_cast<uint8_t *>(std::max(q, reinterpret_cast<uintptr_t>(lo)));
        __atomic_fetch_or(b, (uint8_t)0, __ATOMIC_RELAXED);
    }
}

// One frame's copy work: slabs claimed through an atomic counter by the helper threads and, once its stream has drained, by the
// rendering thread itself.
struct CopyJob {
    const uint8_t *stage = nullptr;  // the shard's compact image
    uint8_t *dst = nullptr;          // compact: the shard's image; framed: the WHOLE frame
    size_t row_bytes = 0;
    uint32_t rows = 0, slab_rows = 0, nslabs = 0, blocks_per_slab = 0;
    volatile uint32_t *done = nullptr;
    bool framed = false;
    uint32_t band_rows = 1, shard_index = 0, shard_count = 1;
    std::atomic<uint32_t> next{0};
    std::atomic<bool> abort{false};
    std::atomic<int> running{0};
    // first touch ahead of the copies: pieces of the destination (512 KB of a compact image, one band of a framed one) are
    // populated by whichever thread is free while the first slabs are still being iterated
    uint32_t npieces = 0;
    std::atomic<uint32_t> next_piece{0};
    static constexpr size_t kPieceBytes = (size_t)512 << 10;

    void plan_populate()
    {
        if (!tuning().populate || !dst || rows == 0) return;
        const size_t bytes = (size_t)rows * row_bytes;
        npieces = framed ? (rows + band_rows - 1) / band_rows : (uint32_t)((bytes + kPieceBytes - 1) / kPieceBytes);
    }
    void populate_piece(uint32_t i) const
    {
        uint8_t *lo, *hi;
        if (framed) {
            const uint32_t r0 = i * band_rows, r1 = std::min(rows, r0 + band_rows);
            lo = dst + ((size_t)i * shard_count + shard_index) * band_rows * row_bytes;
            hi = lo + (size_t)(r1 - r0) * row_bytes;
        } else {
            const size_t bytes = (size_t)rows * row_bytes;
            lo = dst + (size_t)i * kPieceBytes;
            hi = dst + std::min(bytes, (size_t)(i + 1) * kPieceBytes);
        }
        populate_range(lo, hi);
    }

    // A slab is copied in kSlabParts parts, each claimed on its own: what stays exposed at the end of a frame is the copy of the
    // LAST slab, and four threads finish it in a quarter of the time (C2: 768 KB, ~77 us for one thread).
    static constexpr uint32_t kSlabParts = 4;
    void copy_slab(uint32_t s, uint32_t part) const
    {
        const uint32_t s0 = s * slab_rows, s1 = std::min(rows, s0 + slab_rows);
        const uint32_t per = (s1 - s0 + kSlabParts - 1) / kSlabParts;
        const uint32_t r0 = std::min(s1, s0 + part * per), r1 = std::min(s1, r0 + per);
        if (r0 >= r1) return;
        if (!framed) {
            memcpy(dst + (size_t)r0 * row_bytes, stage + (size_t)r0 * row_bytes, (size_t)(r1 - r0) * row_bytes);
            return;
        }
        for (uint32_t lr = r0; lr < r1;) {  // runs of rows inside one band
            const uint32_t band = lr / band_rows, in_band = lr - band * band_rows;
            const uint32_t run = std::min(r1 - lr, band_rows - in_band);
            const size_t y = ((size_t)band * shard_count + shard_index) * band_rows + in_band;
            memcpy(dst + y * row_bytes, stage + (size_t)lr * row_bytes, (size_t)run * row_bytes);
            lr += run;
        }
    }
    // Claims slabs until none is left (or the frame was abandoned).  Every participant was counted in `running` beforehand.
    // `watch` (the rendering thread only): the stream whose completion implies every flag has been raised.  The flags of the
    // sliced path come from host functions in stream order, and CUDA does not run those once the stream or the context is in an
    // error state (a kernel fault, a failed copy) — without a look at the stream every participant would spin for ever, the
    // rendering thread with the device's mutex held.  So while it waits, the watcher polls the stream now and then: an error, or
    // a drained stream with the flag still down, abandons the frame (watch_error says why).
    cudaError_t watch_error = cudaSuccess;
    void run(cudaStream_t watch = nullptr)
    {
        for (;;) {
            const uint32_t i = next_piece.load(std::memory_order_relaxed) < npieces ? next_piece.fetch_add(1, std::memory_order_relaxed) : npieces;
            if (i >= npieces || abort.load(std::memory_order_relaxed)) break;
            populate_piece(i);
        }
        for (;;) {
            const uint32_t item = next.fetch_add(1, std::memory_order_relaxed);
            const uint32_t s = item / kSlabParts;
            if (s >= nslabs) break;
            uint32_t spins = 0;
            while (done[s] == 0u) {
                if (abort.load(std::memory_order_relaxed)) goto out;
                if (++spins < 4096u) {
#if defined(__x86_64__)
                    __builtin_ia32_pause();
#endif
                } else {
                    std::this_thread::yield();
                    if (watch != nullptr && (spins & 63u) == 0u) {
                        const cudaError_t q = cudaStreamQuery(watch);
                        if (q != cudaErrorNotReady && done[s] == 0u) {
                            watch_error = q != cudaSuccess ? q : cudaErrorUnknown;  // drained, yet the callback never ran
                            abort.store(true);
                        }
                    }
                }
            }
            std::atomic_thread_fence(std::memory_order_acquire);  // the slab's pixels were written before its count
            copy_slab(s, item % kSlabParts);
        }
    out:
        running.fetch_sub(1, std::memory_order_release);
    }
    void wait_idle() const
    {
        while (running.load(std::memory_order_acquire) != 0) std::this_thread::yield();
    }
};

// A few host threads shared by all renders of the process (several devices may stage at once); parked on a condition variable.
class CopyPool {
    std::mutex mu_;
    std::condition_variable cv_;
    std::vector<std::thread> threads_;
    std::vector<std::shared_ptr<CopyJob>> queue_;
    bool stop_ = false;

    void loop()
    {
        for (;;) {
            std::shared_ptr<CopyJob> job;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || !queue_.empty(); });
                if (stop_) return;
                job = queue_.back();
                queue_.pop_back();
            }
            job->run();
        }
    }

public:
    int helpers()
    {
        if (tuning().copy_threads > 0) return std::min(16, tuning().copy_threads);
        // fresh 6 MB / 27 MB destinations (C3 / C2) on the 16-core box: 3.27 / 5.47 ms with 3 helpers, 2.35 / 5.08 ms with 6
        // (profiles/r2_pageable_populate.txt)
        const unsigned hc = std::thread::hardware_concurrency();
        return hc >= 16 ? 6 : (hc >= 8 ? 3 : (hc >= 4 ? 2 : 1));
    }
    // `n` helper threads will call job->run(); the caller has added them to job->running
    void post(const std::shared_ptr<CopyJob> &job, int n)
    {
        std::lock_guard<std::mutex> lk(mu_);
        for (int i = 0; i < n; ++i) queue_.push_back(job);
        // a helper stays with its job until the last slab is claimed, so frames staged at the same time (one per device of an
        // in-process multi-GPU render) each need their own: grow up to half the cores
        const int cap = std::max(n, std::min(32, (int)std::thread::hardware_concurrency() / 2));
        const int want = std::min((int)queue_.size(), cap);
        while ((int)threads_.size() < want) threads_.emplace_back([this] { loop(); });
        cv_.notify_all();
    }
    void shutdown()
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
            cv_.notify_all();
        }
        for (auto &t : threads_) t.join();
        threads_.clear();
        queue_.clear();
        stop_ = false;
    }
    ~CopyPool() { shutdown(); }  // a process that exits without frb_shutdown(): the helpers are parked, joining is immediate
};
CopyPool g_copy;
