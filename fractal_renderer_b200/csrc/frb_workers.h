// frb_workers.h — one persistent host worker thread per device (frb_render with device = -1, frb_render_frames).  Included by
// frb_api.cu inside its anonymous namespace; not a stand-alone header.
#pragma once

// ---- one host worker thread per device -------------------------------------------------------------------------------------
// frb_render(device = -1) drives every GPU of the box from one call.  Enqueued from the calling thread alone, device g's kernels
// start ~35 us x g late (plan, ~10 launches and event records per device): at 8 GPUs the last shard of a 5 ms frame started
// 0.28 ms after the first.  Each device therefore has a persistent worker that plans, enqueues, waits and harvests its own
// shard; the caller runs device 0's share itself and joins the rest.  Workers are created on first use, parked on a condition
// variable in between (after a short spin), and joined by frb_shutdown.
struct DeviceWorker {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<void()> job;
    // `state` is the fast path: 0 = idle, 1 = job posted, 2 = running.  A worker that has just finished a job spins on it for a
    // short while before parking on the condition variable, so that during a burst of frames (an animation, an interactive
    // drag, a benchmark loop) the next shard starts within a microsecond instead of a futex wake-up (~20-50 us, comparable
    // to the launch sequence of a whole shard).
    std::atomic<int> state{0};
    bool stop = false;
    static constexpr int kSpinUs = 300;

    void loop()
    {
        for (;;) {
            // spin, then park
            const auto t0 = std::chrono::steady_clock::now();
            bool have = false;
            while (!have) {
                if (state.load(std::memory_order_acquire) == 1) {
                    have = true;
                    break;
                }
                if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(kSpinUs)) break;
#if defined(__x86_64__)
                __builtin_ia32_pause();
#endif
            }
            if (!have) {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [&] { return state.load(std::memory_order_acquire) == 1 || stop; });
                if (stop) return;
            }
            state.store(2, std::memory_order_relaxed);
            job();
            {
                std::lock_guard<std::mutex> lk(m);  // pairs with wait(): the store is not missed between its check and its sleep
                state.store(0, std::memory_order_release);
            }
            cv.notify_all();
        }
    }
    // one job at a time per worker; the previous one has been waited for (frb_render joins every worker before it returns)
    void submit(std::function<void()> j)
    {
        wait();
        job = std::move(j);
        {
            std::lock_guard<std::mutex> lk(m);
            state.store(1, std::memory_order_release);
        }
        cv.notify_all();
    }
    void wait()
    {
        for (int i = 0; i < 2000; ++i) {
            if (state.load(std::memory_order_acquire) == 0) return;
#if defined(__x86_64__)
            __builtin_ia32_pause();
#endif
        }
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return state.load(std::memory_order_acquire) == 0; });
    }
};
DeviceWorker *g_workers[kMaxDevices] = {};

DeviceWorker *get_worker(int dev)
{
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_workers[dev]) {
        DeviceWorker *w = new DeviceWorker();
        w->th = std::thread([w] { w->loop(); });
        g_workers[dev] = w;
    }
    return g_workers[dev];
}

void stop_workers_locked()
{
    for (int d = 0; d < kMaxDevices; ++d) {
        DeviceWorker *w = g_workers[d];
        if (!w) continue;
        {
            std::lock_guard<std::mutex> lk(w->m);
            w->stop = true;
            w->cv.notify_all();
        }
        w->th.join();
        delete w;
        g_workers[d] = nullptr;
    }
}
