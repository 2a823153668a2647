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
    // one job at a time per worker, from one submitter at a time: callers hold g_dispatch_mu (run_parts_on_workers) from their
    // first submit to their last wait, so the previous job has been waited for by whoever posted it
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

// Runs part(0) on the calling thread and part(1) .. part(G-1) on host threads of their own, and returns when all are done.
// The persistent workers take ONE job at a time and a DeviceWorker has no notion of whose job it runs, so the set of them
// belongs to one multi-device call at a time (g_dispatch_mu, held from the first submit to the last wait): two threads inside
// frb_render(device = -1) / frb_render_frames at once — the GUI's render thread and its screenshot thread, src/gui.rs:60 and
// :325 — used to overwrite each other's `job`, and the loser returned FRB_OK with its bands never rendered.  A caller that
// finds the workers taken does not queue behind them (that could be a whole animation): it starts plain threads for this one
// call (~50 us each), whose shards then interleave with the other call's on the devices' own mutexes, frame by frame.
std::mutex g_dispatch_mu;

template <typename Part>
void run_parts_on_workers(uint32_t G, Part &part)
{
    std::unique_lock<std::mutex> pool(g_dispatch_mu, std::try_to_lock);
    if (pool.owns_lock()) {
        std::vector<DeviceWorker *> workers(G, nullptr);
        for (uint32_t g = 1; g < G; ++g) {
            workers[g] = get_worker((int)g);
            workers[g]->submit([&part, g] { part(g); });
        }
        part(0);
        for (uint32_t g = 1; g < G; ++g) workers[g]->wait();
        return;
    }
    std::vector<std::thread> own;
    own.reserve(G);
    for (uint32_t g = 1; g < G; ++g) own.emplace_back([&part, g] { part(g); });
    part(0);
    for (std::thread &t : own) t.join();
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
