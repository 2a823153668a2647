// frb_server.cu — headless frame server: the reference GUI's render thread without the window.
//
// Mirrors src/gui.rs:37-48 (request_redraw: if a render is in flight only remember that another one is wanted —
// latest state wins), :56-82 (render thread: recv config -> get_image -> RGBA8 ColorImage -> store in the shared slot,
// working = false) and :104-117 (the UI side takes the finished image and re-requests if try_redraw was set).
// The GL texture upload and the egui widgets are out of scope (no graphics stack); the RGBA frame is handed to the
// caller instead.  Frames are produced by frb_render(..., FRB_RGBA32, ...) — or by an injected renderer, which is how
// the coalescing logic is tested on machines without a GPU.
#include "../../include/frb.h"

#include <cuda_runtime.h>

#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <random>
#include <string>
#include <thread>
#include <utility>
#include <vector>

namespace {

// A grow-only frame buffer.  With the library's own renderer the frames are PINNED host memory, so frb_render writes them over
// PCIe while the iterations run (streaming finisher) or copies them at DMA speed instead of through the driver's staging
// buffers; with an injected renderer (tests without a GPU) or if pinning fails it is ordinary memory.
struct FrameBuf {
    uint8_t *p = nullptr;
    size_t cap = 0;
    bool pinned = false;
    FrameBuf() = default;
    FrameBuf(const FrameBuf &) = delete;
    FrameBuf &operator=(const FrameBuf &) = delete;
    ~FrameBuf() { release(); }
    void release()
    {
        if (p) {
            if (pinned) cudaFreeHost(p);
            else free(p);
        }
        p = nullptr;
        cap = 0;
    }
    bool reserve(size_t n, bool want_pinned)
    {
        if (n <= cap) return true;
        release();
        const size_t want = n + n / 8 + 4096;  // head-room: interactive resizes should not re-pin every frame
        if (want_pinned) {
            void *q = nullptr;
            if (cudaHostAlloc(&q, want, cudaHostAllocPortable) == cudaSuccess) {
                p = static_cast<uint8_t *>(q);
                pinned = true;
                cap = want;
                return true;
            }
            cudaGetLastError();
        }
        p = static_cast<uint8_t *>(malloc(want));
        pinned = false;
        cap = p ? want : 0;
        return p != nullptr;
    }
    void swap(FrameBuf &o)
    {
        std::swap(p, o.p);
        std::swap(cap, o.cap);
        std::swap(pinned, o.pinned);
    }
};

struct Server {
    int device = 0;
    FrbServerRenderFn renderer = nullptr;
    void *user = nullptr;

    std::mutex mu;
    std::condition_variable cv_work;   // render thread waits for a request
    std::condition_variable cv_done;   // callers wait for a finished frame
    bool quit = false;
    // gui.rs state
    bool working = false;              // `working: Arc<AtomicBool>`
    bool try_redraw = false;           // `try_redraw: bool`
    FrbConfig latest = {};             // the App's current state.config (what a re-request sends)
    bool have_request = false;         // the mpsc channel holds at most one config here
    FrbConfig request = {};
    // the shared `image: Arc<Mutex<Option<ColorImage>>>`
    FrameBuf image;
    FrameBuf held;                     // the frame last handed out by frb_server_acquire (owned by the caller until the next one)
    bool pin_frames = false;           // the library's own renderer: frames in pinned memory
    uint32_t image_w = 0, image_h = 0;
    bool image_fresh = false;
    uint64_t frames_rendered = 0, requests = 0, coalesced = 0;
    int last_rc = FRB_OK;
    std::string last_error;
    std::thread thread;
};

// What the reference's render thread calls: get_image (src/gui.rs:60), which dispatches on the algorithm (src/lib.rs:254-320).
// Mandelbrot | Julia -> the per-pixel path; BarnsleyFern -> the chaos game with rayon's thread count and a fresh entropy seed.
int default_renderer(const FrbConfig *cfg, uint8_t *rgba, size_t len, void *user)
{
    const int device = (int)(intptr_t)user;
    if (cfg->algo == FRB_ALGO_BARNSLEY_FERN) {
        std::random_device rd;  // SmallRng::from_entropy (src/lib.rs:428)
        const uint64_t seed = ((uint64_t)rd() << 32) ^ (uint64_t)rd();
        return frb_render_fern(cfg, 0u, seed, nullptr, rgba, len, FRB_RGBA32, device < 0 ? 0 : device);
    }
    return frb_render(cfg, rgba, len, FRB_RGBA32, device);
}

void render_loop(Server *s)
{
    FrameBuf frame;
    for (;;) {
        FrbConfig cfg;
        {
            std::unique_lock<std::mutex> lk(s->mu);
            s->cv_work.wait(lk, [&] { return s->quit || s->have_request; });
            if (s->quit) return;  // "Shutting rendering down." (src/gui.rs:84)
            cfg = s->request;
            s->have_request = false;
        }
        const size_t bytes = (size_t)cfg.width * cfg.height * 4;
        int rc = FRB_OK;
        if (!frame.reserve(bytes, s->pin_frames)) {
            rc = FRB_ERR_INVALID;
        } else {
            rc = s->renderer(&cfg, frame.p, bytes, s->user);
        }
        {
            std::lock_guard<std::mutex> lk(s->mu);
            s->last_rc = rc;
            if (rc == FRB_OK) {
                s->image.swap(frame);
                s->image_w = cfg.width;
                s->image_h = cfg.height;
                s->image_fresh = true;
                s->frames_rendered++;
            } else {
                s->last_error = frame.p ? frb_last_error() : "out of host memory for the frame";
            }
            s->working = false;
        }
        s->cv_done.notify_all();  // frame.request_repaint()
    }
}

// App::request_redraw (src/gui.rs:37-48); caller holds s->mu
void request_locked(Server *s)
{
    s->requests++;
    if (s->working) {
        s->try_redraw = true;
        s->coalesced++;
        return;
    }
    s->try_redraw = false;
    s->working = true;
    s->request = s->latest;
    s->have_request = true;
    s->cv_work.notify_one();
}

}  // namespace

extern "C" {

FrbServer *frb_server_create(int device, FrbServerRenderFn renderer, void *user)
{
    Server *s = new (std::nothrow) Server();
    if (!s) return nullptr;
    s->device = device;
    s->renderer = renderer ? renderer : default_renderer;
    s->pin_frames = renderer == nullptr;
    s->user = renderer ? user : (void *)(intptr_t)device;
    s->thread = std::thread(render_loop, s);
    return reinterpret_cast<FrbServer *>(s);
}

void frb_server_destroy(FrbServer *h)
{
    Server *s = reinterpret_cast<Server *>(h);
    if (!s) return;
    {
        std::lock_guard<std::mutex> lk(s->mu);
        s->quit = true;
    }
    s->cv_work.notify_all();
    if (s->thread.joinable()) s->thread.join();
    delete s;
}

int frb_server_request(FrbServer *h, const FrbConfig *cfg)
{
    Server *s = reinterpret_cast<Server *>(h);
    if (!s || !cfg) return FRB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(s->mu);
    s->latest = *cfg;
    request_locked(s);
    return FRB_OK;
}

// Waits (up to timeout_ms; < 0 = for ever) for a fresh frame or an error.  Caller holds `lk`.  Returns 1 (fresh frame in the
// slot), 0 (none) or the render error, which it clears.
static int wait_frame(Server *s, std::unique_lock<std::mutex> &lk, int timeout_ms)
{
    if (timeout_ms != 0 && !s->image_fresh && s->last_rc == FRB_OK) {
        auto ready = [&] { return s->image_fresh || s->last_rc != FRB_OK || (!s->working && !s->have_request); };
        if (timeout_ms < 0)
            s->cv_done.wait(lk, ready);
        else
            s->cv_done.wait_for(lk, std::chrono::milliseconds(timeout_ms), ready);
    }
    if (s->last_rc != FRB_OK) {
        const int rc = s->last_rc;
        s->last_rc = FRB_OK;
        return rc;
    }
    return s->image_fresh ? 1 : 0;
}

int frb_server_poll(FrbServer *h, uint8_t *rgba, size_t len, uint32_t *width, uint32_t *height, int timeout_ms)
{
    Server *s = reinterpret_cast<Server *>(h);
    if (!s) return FRB_ERR_INVALID;
    std::unique_lock<std::mutex> lk(s->mu);
    const int st = wait_frame(s, lk, timeout_ms);
    if (st <= 0) return st;
    // `app.image.lock().unwrap().take()` (src/gui.rs:110) and the re-request of :115-117
    const size_t bytes = (size_t)s->image_w * s->image_h * 4;
    if (width) *width = s->image_w;
    if (height) *height = s->image_h;
    if (!rgba) return 1;  // peek: report the size, leave the frame in the slot
    if (len < bytes) return FRB_ERR_INVALID;
    memcpy(rgba, s->image.p, bytes);
    s->image_fresh = false;
    if (s->try_redraw) request_locked(s);
    return 1;
}

int frb_server_acquire(FrbServer *h, const uint8_t **rgba, uint32_t *width, uint32_t *height, int timeout_ms)
{
    Server *s = reinterpret_cast<Server *>(h);
    if (!s || !rgba) return FRB_ERR_INVALID;
    std::unique_lock<std::mutex> lk(s->mu);
    const int st = wait_frame(s, lk, timeout_ms);
    if (st <= 0) return st;
    // the frame moves out of the slot; the buffer handed out last time goes back into circulation
    s->held.swap(s->image);
    s->image_fresh = false;
    *rgba = s->held.p;
    if (width) *width = s->image_w;
    if (height) *height = s->image_h;
    if (s->try_redraw) request_locked(s);
    return 1;
}

int frb_server_stats(FrbServer *h, FrbServerStats *out)
{
    Server *s = reinterpret_cast<Server *>(h);
    if (!s || !out) return FRB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(s->mu);
    out->requests = s->requests;
    out->coalesced = s->coalesced;
    out->frames_rendered = s->frames_rendered;
    out->working = s->working ? 1u : 0u;
    out->try_redraw = s->try_redraw ? 1u : 0u;
    return FRB_OK;
}

const char *frb_server_last_error(FrbServer *h)
{
    Server *s = reinterpret_cast<Server *>(h);
    if (!s) return "";
    std::lock_guard<std::mutex> lk(s->mu);
    return s->last_error.c_str();
}

}  // extern "C"
