// frb_api.cu — the C ABI of include/frb.h: device contexts, buffer pools, launches, copies, statistics.
//
// Replaces the Mandelbrot|Julia arm of the reference's get_image (src/lib.rs:253-270).  No CPU fallback exists:
// without a usable CUDA device every compute entry point fails with FRB_ERR_NO_DEVICE.
#include "../../include/frb.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "frb_kernels.cuh"
#include "frb_fern.cuh"

#include <thread>

#include <sys/mman.h>

namespace {

using frb::RenderParams;

constexpr int kMaxDevices = 32;
constexpr int kSlots = 64;               // in-flight async renders per device whose statistics stay readable
constexpr uint32_t kDefaultChunk = 64;   // pixels per queue chunk
constexpr uint32_t kDefaultBandRows = 16;
constexpr uint32_t kMinDeferIterations = 32;  // the deferred kernel needs a few blocks of iterations to be worth its resolve pass
constexpr uint32_t kMaxDeferIterations = 1u << 30;  // its budget counter shares a word with the escape filter (frb_deferred.cuh)
constexpr uint32_t kMaxStageSlabs = 16384;  // upper bound on the finisher slabs of a frame that goes through pinned staging
constexpr uint32_t kMaxIterWarps = 16384;  // upper bound on resident warps of one iterate launch (SMs x CTAs/SM x 8)

// ---- tuning -----------------------------------------------------------------------------------------------------------------
// Every choice the library makes per frame (block shape of the deferred kernel, chunk size, slab counts, which host path) has
// its default here.  A product build (no -DFRB_TUNING) reads nothing else: no environment variable, no hidden switch, the
// defaults are compile-time constants.  The tuning build (`make tune` -> libfrb_tune.so, used by the sweeps under scripts/ and by
// two tests that force rarely taken paths) makes them settable through frb_set_tuning().
// Tile kernel shapes (frb_tile.cuh): 8 x 4 pixels per warp for images in device memory (best lane utilisation: C1 0.930 against
// 0.914 for 16 x 2 and 0.884 for 32 x 1), 32 x 1 when the kernel's stores go straight to pinned host memory: a tile row is then
// 96 contiguous bytes on the PCIe link instead of 24 (C1 end to end: 0.187 ms against 0.269; profiles/r2_tile_shapes.txt)
constexpr int kTileDefaultW = 8, kTileHostW = 32, kTileDefaultWpc = 4;
struct Tuning {
    int defer_b = 0;            // deferred kernel: iterations per block (0 = by iteration cap)
    int defer_nb = 0;           //                  blocks per trip      (0 = by iteration cap)
    int defer_shape = -1;       //                  index into kDeferShapes (-1 = default)
    int chunk = 0;              // pixels per queue chunk (0 = by iteration cap)
    int fast_shade = 1;         // 0 = colour through libdevice log2 and the division sequence
    int tile_max = 384;         // iteration caps up to this take the lock-step tile kernel
    int stream = 1;             // 0 = never use the streaming finisher
    int stream_gen_polls = 20000;  // finisher: bound on the wait for the frame to start (~1 us per poll); 0 forces the fallback
    int stream_rows = 0;        // finisher: rows per slab (0 = ~256 Kpx)
    int stream_ctas = 1;        // finisher CTAs per SM
    int slabs = 0;              // sliced host path: number of row slabs (0 = by frame size)
    int no_slab = 0;            // 1 = never slice
    int stage = 1;              // 0 = pageable destinations take the sliced copy path instead of the pinned staging pipeline
    int copy_threads = 0;       // staging pipeline: host copy threads (0 = default)
    int end_chunk = 0;          // deferred kernel: pixels per chunk in the end phase of the queue (0 = default)
    int hugepage = 1;           // madvise(MADV_HUGEPAGE) on large pageable destinations
    int tile_direct = 1;        // tile kernel writes a pinned compact destination directly (0 = slabs + copies)
    int pool = 1;               // deferred kernel: 0 = no orbit pool (drained warps keep their lanes to the end)
    int pool_max = 31;          //                  a drained warp with at most this many live lanes donates them (<= 31)
    int populate = 1;           // staging pipeline: first-touch the destination ahead of the copies
    int tile_smem_kb = 0;       // tile kernel: unused dynamic shared memory per CTA, to cap the CTAs resident per SM (experiments)
    int tile_w = 0;                 // tile kernel: pixels per tile row (8 x 4, 16 x 2 or 32 x 1 pixels per warp; 0 = by destination)
    int tile_wpc = kTileDefaultWpc; //              warps per CTA
};
#ifdef FRB_TUNING
Tuning g_tuning;
inline const Tuning &tuning() { return g_tuning; }
#else
constexpr Tuning kTuning = {};
inline const Tuning &tuning() { return kTuning; }
#endif

thread_local std::string tl_error;
thread_local int tl_code = 0;
thread_local FrbStats tl_stats = {};
thread_local uint32_t tl_variant = FRB_KERNEL_AUTO;

int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    tl_error = buf;
    tl_code = code;
    return code;
}

#define CU(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return fail(FRB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct StatSlot {
    unsigned long long iterations;
    unsigned long long lane_iterations;
    unsigned long long stream_status;  // low word != 0: the streaming finisher gave up waiting (frb_stream.cuh)
    unsigned long long reserved[9];    // device words 3..11 (trace stamps, election, work counter, generation)
    unsigned long long stream_started; // device word 12: finisher CTAs that saw their frame start (0 = nothing was shaded)
    unsigned long long reserved2[3];   // device words 13..15
    unsigned long long tile_part[2 * frb::kTileStatsPairs];  // the tile kernel's spread (iterations, lane slots) counters
};
static_assert(offsetof(StatSlot, tile_part) == frb::kTileStatsWord * sizeof(unsigned long long), "StatSlot mirrors the device words");
constexpr size_t kStatsBytes = 1024;  // device statistics block
static_assert(sizeof(StatSlot) <= kStatsBytes, "device statistics block");

struct Slot {
    cudaEvent_t ev_start = nullptr, ev_k0 = nullptr, ev_iter = nullptr, ev_shade = nullptr, ev_end = nullptr;
    bool have_k0 = false;  // ev_k0 (after the table kernel) was recorded for this frame: iterate_ms is the iterate kernel alone
    StatSlot *host = nullptr;  // pinned
    uint64_t pixels = 0;
    uint32_t launches = 0;
    uint32_t variant = 0;
    bool shaded = false;
    bool stream_fallback = false;  // the streamed frame was re-shaded by the ordinary kernel (finisher never started or gave up)
    uint64_t tile_key = 0;  // non-zero: the tile kernel was the library's choice for this frame (see tile_key_of)
};

struct DeviceCtx {
    int dev = -1;
    bool ready = false;
    int sm_count = 0;
    int iter_blocks_per_sm[2] = {0, 0};  // resident CTAs per SM of the dense kernels: FORM 6, FORM 7
    cudaStream_t stream = nullptr;       // the library's own stream for this device
    // tile-kernel feedback: the scene (tile_key_of) whose last tile-kernel frame ran with poor lane utilisation, and for how
    // many more frames the library keeps the lane-refill kernel for it before trying the tile kernel again
    std::vector<const void *> loaded_kernels;  // preload_kernel(): kernels whose code is known to be on the device
    // callers blocked on `mu` in a blocking render: frb_render_frames gives way to them between two frames (a std::mutex is
    // not fair: a thread that unlocks and locks again at once can starve a waiter for the whole animation)
    std::atomic<int> waiting{0};
    std::atomic<uint64_t> tile_bad_key{0};
    std::atomic<uint32_t> tile_bad_frames{0};
    cudaStream_t aux[2] = {nullptr, nullptr};  // slab streams of the host-output path
    cudaEvent_t ev_fork = nullptr, ev_join[2] = {nullptr, nullptr}, ev_slab[16] = {};
    // Two sets of per-render device buffers: blocking and async renders use set 0; frb_render_frames alternates sets
    // (and streams) so that frame k+1's iterations fill the tail of frame k and hide its copy-out.
    struct BufSet {
        cudaEvent_t ev_guard = nullptr;  // completion of the last render that used this set
        cudaStream_t last_stream = nullptr;
        uint64_t last_use = 0;           // DeviceCtx::use_tick of that render (frb_render_device_async gives a new stream the older set)
        bool have_guard = false;
        double *tab = nullptr;           // grow-only
        size_t tab_cap = 0;              // doubles
        uint32_t *iters = nullptr;
        size_t iters_cap = 0;            // elements
        double2 *zfinal = nullptr;
        size_t z_cap = 0;
        uint8_t *out = nullptr;
        size_t out_cap = 0;              // bytes
        unsigned int *queue = nullptr;   // 16 chunk counters
        uint32_t *wm = nullptr;          // kMaxIterWarps watermarks of the deferred kernel's warps (streaming finisher)
        unsigned long long *stats = nullptr;
    } sets[2];
    int cur_set = 0;
    uint64_t use_tick = 0;
    BufSet &b() { return sets[cur_set]; }
    // Barnsley fern (frb_render_fern): per-virtual-thread hit counts and the colour table, grow-only
    uint32_t *fern_counts = nullptr;
    size_t fern_counts_cap = 0;
    uint32_t *fern_lut = nullptr;
    uint8_t *fern_out = nullptr;
    size_t fern_out_cap = 0;
    cudaEvent_t ev_fern[2] = {nullptr, nullptr};
    std::mutex mu;
    Slot slots[kSlots];
    uint32_t next_slot = 0;
    uint32_t stream_gen = 0;             // generation counter of streamed frames (never 0 in a frame)
};

std::mutex g_mu;
DeviceCtx *g_ctx[kMaxDevices] = {};
int g_device_count = -1;
uint32_t g_device_mask = 0;  // devices chosen by frb_init (0 = not initialised)
std::vector<int> g_device_list;  // frb_init_list: the shards of a device = -1 render go to these, in this order (may repeat)

int device_count_locked()
{
    if (g_device_count < 0) {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess) {
            cudaGetLastError();
            n = 0;
        }
        g_device_count = std::min(n, kMaxDevices);
    }
    return g_device_count;
}

int ctx_init(DeviceCtx *c, int dev)
{
    CU(cudaSetDevice(dev));
    c->dev = dev;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10)
        return fail(FRB_ERR_NO_DEVICE, "device %d is sm_%d%d; this library carries sm_100a code only", dev, prop.major, prop.minor);
    c->sm_count = prop.multiProcessorCount;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (auto &bs : c->sets) {
        CU(cudaEventCreateWithFlags(&bs.ev_guard, cudaEventDisableTiming));
        CU(cudaMalloc(&bs.queue, 256));
        CU(cudaMalloc(&bs.stats, kStatsBytes));
        CU(cudaMemset(bs.stats, 0, kStatsBytes));
        CU(cudaMalloc(&bs.wm, (kMaxIterWarps + kMaxStageSlabs) * sizeof(uint32_t)));  // watermarks, then the finisher's slab counters
    }
    CU(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    for (auto &e : c->ev_slab) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (int i = 0; i < 2; ++i) {
        CU(cudaStreamCreateWithFlags(&c->aux[i], cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&c->ev_join[i], cudaEventDisableTiming));
    }
    for (auto &s : c->slots) {
        CU(cudaEventCreate(&s.ev_start));
        CU(cudaEventCreate(&s.ev_iter));
        CU(cudaEventCreate(&s.ev_k0));
        CU(cudaEventCreate(&s.ev_shade));
        CU(cudaEventCreate(&s.ev_end));
        CU(cudaMallocHost(&s.host, sizeof(StatSlot)));
        memset(s.host, 0, sizeof(StatSlot));
    }
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->iter_blocks_per_sm[0], frb::frb_iterate_stream_kernel<6>,
                                                     frb::kIterateThreads, 0));
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c->iter_blocks_per_sm[1], frb::frb_iterate_stream_kernel<7>,
                                                     frb::kIterateThreads, 0));
    if (c->iter_blocks_per_sm[0] < 1 || c->iter_blocks_per_sm[1] < 1)
        return fail(FRB_ERR_CUDA, "iterate kernel does not fit on device %d", dev);
    c->ready = true;
    return FRB_OK;
}

void ctx_destroy(DeviceCtx *c)
{
    if (!c) return;
    if (c->dev >= 0 && cudaSetDevice(c->dev) == cudaSuccess) {
        cudaDeviceSynchronize();
        for (auto &bs : c->sets) {
            cudaFree(bs.tab);
            cudaFree(bs.iters);
            cudaFree(bs.zfinal);
            cudaFree(bs.out);
            cudaFree(bs.queue);
            cudaFree(bs.stats);
            cudaFree(bs.wm);
            if (bs.ev_guard) cudaEventDestroy(bs.ev_guard);
        }
        for (auto &s : c->slots) {
            if (s.ev_start) cudaEventDestroy(s.ev_start);
            if (s.ev_iter) cudaEventDestroy(s.ev_iter);
            if (s.ev_k0) cudaEventDestroy(s.ev_k0);
            if (s.ev_shade) cudaEventDestroy(s.ev_shade);
            if (s.ev_end) cudaEventDestroy(s.ev_end);
            if (s.host) cudaFreeHost(s.host);
        }
        cudaFree(c->fern_counts);
        cudaFree(c->fern_lut);
        cudaFree(c->fern_out);
        for (auto &e : c->ev_fern)
            if (e) cudaEventDestroy(e);
        if (c->ev_fork) cudaEventDestroy(c->ev_fork);
        for (auto &e : c->ev_slab)
            if (e) cudaEventDestroy(e);
        for (int i = 0; i < 2; ++i) {
            if (c->ev_join[i]) cudaEventDestroy(c->ev_join[i]);
            if (c->aux[i]) cudaStreamDestroy(c->aux[i]);
        }
        if (c->stream) cudaStreamDestroy(c->stream);
    }
    delete c;
}

// Returns the (lazily created) context of `dev`, or nullptr with the error set.
DeviceCtx *get_ctx(int dev)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const int n = device_count_locked();
    if (n == 0) {
        fail(FRB_ERR_NO_DEVICE, "no usable CUDA device (cudaGetDeviceCount found none); this library has no CPU path");
        return nullptr;
    }
    if (dev < 0 || dev >= n) {
        fail(FRB_ERR_INVALID, "device %d out of range [0, %d)", dev, n);
        return nullptr;
    }
    if (!g_ctx[dev]) {
        DeviceCtx *c = new DeviceCtx();
        if (ctx_init(c, dev) != FRB_OK) {
            const std::string keep = tl_error;
            const int code = tl_code;
            ctx_destroy(c);
            tl_error = keep;
            tl_code = code;
            return nullptr;
        }
        g_ctx[dev] = c;
    }
    return g_ctx[dev];
}

template <typename T>
int ensure(T *&ptr, size_t &cap, size_t need)
{
    if (need <= cap) return FRB_OK;
    if (ptr) CU(cudaFree(ptr));
    ptr = nullptr;
    cap = 0;
    const size_t want = need + need / 8 + 1024;  // a little head-room so interactive resizes do not thrash
    CU(cudaMalloc(&ptr, want * sizeof(T)));
    cap = want;
    return FRB_OK;
}

uint64_t double_bits(double v)
{
    uint64_t b;
    memcpy(&b, &v, sizeof b);
    return b;
}

FrbShard whole_frame()
{
    FrbShard s = {0u, 1u, 0u, 0u};
    return s;
}

bool shard_valid(const FrbShard &s) { return s.count >= 1 && s.index < s.count; }

uint32_t shard_band_rows(uint32_t height, const FrbShard &s)
{
    if (s.band_rows == 0) return std::max(height, 1u);
    return s.band_rows;
}

uint32_t shard_rows(uint32_t height, const FrbShard &s)
{
    if (!shard_valid(s)) return 0;
    const uint64_t br = shard_band_rows(height, s);
    const uint64_t nbands = (height + br - 1) / br;  // bands in the image
    uint64_t rows = 0;
    if (nbands > s.index) {
        const uint64_t mine = (nbands - s.index + s.count - 1) / s.count;  // bands k = index, index+count, ...
        rows = mine * br;
        const uint64_t last_band = s.index + (mine - 1) * s.count;
        const uint64_t last_end = (last_band + 1) * br;
        if (last_end > height) rows -= (last_end - height);
    }
    return (uint32_t)rows;
}

uint32_t shard_row_to_y(uint32_t height, const FrbShard &s, uint32_t lr)
{
    const uint32_t br = shard_band_rows(height, s);
    const uint32_t band = lr / br;
    return (band * s.count + s.index) * br + (lr - band * br);
}

// coord_to_space on the host with the same operation order as the device table kernel (used only to decide
// which iteration form is provably exact; never to produce pixels).
double host_coord(double coord, double max, double offset, double pos, double scale)
{
    volatile double a = coord / max;
    volatile double b = a - offset;
    volatile double c = b / scale;
    volatile double d = c + pos;
    return d;
}

// FORM 6 (p = x*y; nim = fma(p, 2, c.im)) differs from the reference's (2x*y) + c.im only when x*y is subnormal
// or 2*x*y overflows.  It is selected only when neither can alter a result bit (see frb_kernels.cuh).
bool form6_is_exact(const FrbConfig &cfg)
{
    const double S = cfg.limit * cfg.limit;
    if (!(S <= 1e300)) return false;  // also rejects NaN and inf
    const double tiny = 0x1p-900, huge = 1e150;
    const double wf = (double)cfg.width, hf = (double)cfg.height;
    // start points must not overflow x*y on the very first product
    const double re_a = host_coord(0.0, hf, (wf / hf) / 2.0, cfg.pos.re, cfg.scale.re);
    const double re_b = host_coord((double)(cfg.width ? cfg.width - 1 : 0), hf, (wf / hf) / 2.0, cfg.pos.re, cfg.scale.re);
    const double im_a = host_coord(0.0, hf, 0.5, cfg.pos.im, cfg.scale.im);
    const double im_b = host_coord((double)(cfg.height ? cfg.height - 1 : 0), hf, 0.5, cfg.pos.im, cfg.scale.im);
    const double ext[4] = {re_a, re_b, im_a, im_b};
    for (double v : ext)
        if (!(std::fabs(v) <= huge)) return false;
    if (cfg.algo == FRB_ALGO_JULIA) {
        if (!(std::fabs(cfg.julia_set.re) <= huge) || !(std::fabs(cfg.julia_set.im) <= huge)) return false;
        return std::fabs(cfg.julia_set.im) >= tiny;  // c.im == 0 lets im(z) decay through the subnormals: use FORM 7
    }
    // Mandelbrot: c.im = im0(y).  Exactly 0 is fine (im(z) stays 0); a non-zero |im0| < 2^-900 is not.
    // im0 is monotone in y, so the smallest non-zero magnitude sits next to the sign change.
    if ((im_a > 0.0) == (im_b > 0.0) && im_a != 0.0 && im_b != 0.0)
        return std::min(std::fabs(im_a), std::fabs(im_b)) >= tiny;
    for (uint32_t y = 0; y < cfg.height; ++y) {
        const double v = std::fabs(host_coord((double)y, hf, 0.5, cfg.pos.im, cfg.scale.im));
        if (v != 0.0 && !(v >= tiny)) return false;
    }
    return true;
}

// The deferred kernel tests for escape once per block of B iterations and its resolve pass re-runs at most B + 1 steps after a
// stop by the filter.  Both are only sound when an orbit that fired the filter provably escapes at once and keeps growing
// (frb_deferred.cuh): the filter compares high words, i.e. fires for max(re^2, im^2) >= (S/2)(1 - 2^-20), so the next step has
// |z| >= (R^2/2)(1 - 2^-20) - |c|, and that must exceed R with room to spare: R = |limit| in [8, 2^20] and |c| <= R^2/2 - 3R for
// every pixel give |z| >= 3R - R^2 2^-21 >= 2.5 R.  (Above R ~ 4.7e6 the truncation slack R^2 2^-21 eats the margin: a near-miss
// stop need not escape within the resolve bound.  Larger limits take the dense kernel, which tests every iteration.)
bool deferred_is_safe(const FrbConfig &cfg)
{
    if (!form6_is_exact(cfg)) return false;
    if (cfg.iterations < kMinDeferIterations) return false;  // too short to amortise a block
    if (cfg.iterations >= kMaxDeferIterations) return false;
    const double R = std::fabs(cfg.limit);
    if (!(R >= 8.0) || !(R <= 1048576.0)) return false;
    double cmax;
    if (cfg.algo == FRB_ALGO_JULIA) {
        cmax = std::hypot(cfg.julia_set.re, cfg.julia_set.im);
    } else {
        const double wf = (double)cfg.width, hf = (double)cfg.height;
        const double re_a = host_coord(0.0, hf, (wf / hf) / 2.0, cfg.pos.re, cfg.scale.re);
        const double re_b = host_coord((double)(cfg.width ? cfg.width - 1 : 0), hf, (wf / hf) / 2.0, cfg.pos.re, cfg.scale.re);
        const double im_a = host_coord(0.0, hf, 0.5, cfg.pos.im, cfg.scale.im);
        const double im_b = host_coord((double)(cfg.height ? cfg.height - 1 : 0), hf, 0.5, cfg.pos.im, cfg.scale.im);
        cmax = std::hypot(std::max(std::fabs(re_a), std::fabs(re_b)), std::max(std::fabs(im_a), std::fabs(im_b)));
    }
    return cmax * 1.0000001 <= 0.5 * R * R - 3.0 * R;
}

// CTA shapes of the deferred kernel: threads per CTA x resident CTAs per SM (the tuning build can pick any by index).  Co-resident
// CTAs are served very unevenly by the warp scheduler (frb_deferred.cuh), hence one wide CTA per SM by default.
struct DeferShape {
    int threads, per_sm;
};
constexpr DeferShape kDeferShapes[] = {{256, 3}, {512, 1}, {768, 1}, {1024, 1}, {512, 2}, {256, 2}};
constexpr int kDeferShapeCount = (int)(sizeof(kDeferShapes) / sizeof(kDeferShapes[0]));
constexpr int kDeferDefaultShape = 2;

// Shape of the deferred kernel (FRB_KERNEL_DEFER6): B iterations per block, NB blocks per trip, CTA shape.
// Lanes that stop idle until the trip ends and a stopped pixel costs ~B/2 wasted + ~B/2 resolved iterations, so short
// pixels want short blocks and trips; the iteration cap is the only a-priori hint.  Measured, profiles/r1_tuning_sweeps.txt.
void choose_deferred_shape(const FrbConfig &cfg, int *b, int *nb, int *minb)
{
    *b = cfg.iterations >= 2048u ? 16 : 8;
    *nb = cfg.iterations >= 2048u ? 4 : (cfg.iterations >= 256u ? 4 : 1);
    *minb = kDeferDefaultShape;
    const Tuning &t = tuning();
    if (t.defer_b) *b = t.defer_b;
    if (t.defer_nb) *nb = t.defer_nb;
    if (t.defer_shape >= 0) *minb = t.defer_shape;
    if (*b != 8 && *b != 16 && *b != 32) *b = 16;
    if (*nb != 1 && *nb != 2 && *nb != 4 && *nb != 8) *nb = 4;
    if (*b == 32 && *nb == 8) *nb = 4;
    if (*b == 32 && *nb == 1) *nb = 2;
    if (*b == 16 && *nb == 1) *nb = 2;
    if (*minb < 0 || *minb >= kDeferShapeCount) *minb = kDeferDefaultShape;
}

struct Plan {
    RenderParams p;
    uint32_t variant;
    bool fern;      // algo = BarnsleyFern: the per-pixel path is all black (calc/src/lib.rs:211); no iterate kernel runs
    int tile_form;  // tile kernel: 6- or 7-instruction iteration
    bool tile_host; //              its stores go to pinned host memory (render_blocking: tile_alias)
    uint64_t tile_key;  // non-zero when AUTO chose the tile kernel
    int defer_b, defer_nb, defer_minb;  // deferred kernel: iterations per block, blocks per trip, CTA shape (index into kDeferShapes)
    uint64_t npix;
    size_t out_bytes;
};

// What "the same scene" means for the tile-kernel feedback: algorithm, cap, escape radius and the Julia constant — not the
// view, which changes every frame of a zoom or a drag.  Never 0.
uint64_t tile_key_of(const FrbConfig &cfg)
{
    uint64_t h = 0x9e3779b97f4a7c15ull;
    const uint64_t w[5] = {cfg.algo, cfg.iterations, double_bits(cfg.limit), double_bits(cfg.julia_set.re), double_bits(cfg.julia_set.im)};
    for (uint64_t x : w) {
        h ^= x + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
        h *= 0xff51afd7ed558ccdull;
    }
    return h | 1ull;
}
constexpr uint32_t kTileBackoffFrames = 64;
constexpr double kTileMinUtilisation = 0.7;  // break-even against the refill kernel: ~0.7 at 40 iterations/px, higher above

int make_plan(DeviceCtx &c, const FrbConfig &cfg, const FrbShard &sh, uint32_t format, Plan *plan)
{
    if (!shard_valid(sh)) return fail(FRB_ERR_INVALID, "bad shard: index %u of count %u", sh.index, sh.count);
    if (format != 0 && format != FRB_RGB24 && format != FRB_RGBA32) return fail(FRB_ERR_INVALID, "unknown pixel format %u", format);
    if (cfg.algo > FRB_ALGO_JULIA) return fail(FRB_ERR_INVALID, "unknown algo %u", cfg.algo);
    RenderParams &p = plan->p;
    memset(&p, 0, sizeof p);
    p.width = cfg.width;
    p.height = cfg.height;
    p.rows = shard_rows(cfg.height, sh);
    p.band_rows = shard_band_rows(cfg.height, sh);
    p.band_shift = 32u;
    for (uint32_t k = 0; k < 32u; ++k)
        if (p.band_rows == (1u << k)) p.band_shift = k;
    p.shard_index = sh.index;
    p.shard_count = sh.count;
    p.iterations = cfg.iterations;
    p.algo = cfg.algo;
    // `let squared = limit * limit` (calc/src/lib.rs:246): one IEEE multiply, same on host and device
    {
        volatile double l = cfg.limit;
        volatile double s = l * l;
        p.limit_sq = s;
    }
    p.filter_hi = (uint32_t)(double_bits(p.limit_sq * 0.5) >> 32);
    {
        const uint32_t h = (uint32_t)(double_bits(std::sqrt(p.limit_sq * 0.5)) >> 32);
        p.filter_abs_hi = h ? h - 1u : 0u;
    }
    // 32-pixel segments balance the end of a long-iteration frame better (C2: 5.00 vs 5.09 ms); short-iteration frames
    // (C1, C3) prefer fewer queue transactions (measured, profiles/r1_tuning_sweeps.txt)
    p.chunk = cfg.iterations >= 2048u ? 32u : kDefaultChunk;
    if (tuning().chunk > 0) p.chunk = (uint32_t)tuning().chunk;
    p.chunks_per_row = cfg.width ? (cfg.width + p.chunk - 1) / p.chunk : 0;
    const uint64_t npix = (uint64_t)p.rows * cfg.width;
    const uint64_t chunks = (uint64_t)p.rows * p.chunks_per_row;
    if (npix > 0xfffffff0ull || chunks > 0xfffffff0ull)
        return fail(FRB_ERR_TOO_LARGE, "shard of %u x %u pixels exceeds 2^32 pixels per launch; use more shards", cfg.width, p.rows);
    p.total_chunks = (uint32_t)chunks;
    p.julia_re = cfg.julia_set.re;
    p.julia_im = cfg.julia_set.im;
    p.pos_re = cfg.pos.re;
    p.pos_im = cfg.pos.im;
    p.scale_re = cfg.scale.re;
    p.scale_im = cfg.scale.im;
    p.stable_limit = cfg.stable_limit;
    p.exposure = cfg.exposure;
    p.smooth = cfg.smooth ? 1u : 0u;
    p.inside = cfg.inside ? 1u : 0u;
    p.primary = (uint32_t)cfg.primary_color.r | ((uint32_t)cfg.primary_color.g << 8) | ((uint32_t)cfg.primary_color.b << 16);
    p.secondary = (uint32_t)cfg.secondary_color.r | ((uint32_t)cfg.secondary_color.g << 8) | ((uint32_t)cfg.secondary_color.b << 16);
    p.format = format;
    // colour fast path (shade_pixel): uniform conversions done once here (the tuning build can switch it off: a test compares
    // it with libdevice's log2 and the division sequence)
    p.iterations_d = (double)cfg.iterations;
    p.inv_iterations = cfg.iterations ? 1.0 / p.iterations_d : 0.0;
    p.prim_r = (double)cfg.primary_color.r;
    p.prim_g = (double)cfg.primary_color.g;
    p.prim_b = (double)cfg.primary_color.b;
    p.fast_shade = (cfg.iterations >= 1u && std::isfinite(cfg.exposure) && std::fabs(cfg.exposure) < 1e290) ? 1u : 0u;
    if (!tuning().fast_shade) p.fast_shade = 0u;
    plan->npix = npix;
    plan->out_bytes = (size_t)npix * format;
    uint32_t v = tl_variant;
    // A forced variant is a request: outside its proven domain it is downgraded to the next kernel that is exact
    // (DEFER6 -> DENSE6 -> STREAM7); FrbStats.kernel_variant reports what actually ran.
    // short orbits, image wanted: the lock-step tile kernel (per-pixel costs of the refill kernels outweigh their lane
    // utilisation there).  It has no raw-result output (format 0 falls through), and its grid puts 8 rows on blockIdx.y.
    const uint32_t tile_max = (uint32_t)std::min(65535, std::max(0, tuning().tile_max));  // its statistics are 32-bit sums per CTA
    const bool tile_fits = p.rows <= 65535u;  // grid.y carries the tile rows
    plan->fern = cfg.algo == FRB_ALGO_BARNSLEY_FERN;  // get_recursive_pixel returns BLACK for it: nothing to iterate
    if (v == FRB_KERNEL_TILE && (format == 0 || plan->fern || !tile_fits || cfg.iterations > 65535u)) v = FRB_KERNEL_AUTO;
    plan->tile_key = 0;
    if (v == FRB_KERNEL_AUTO && format != 0 && cfg.iterations <= tile_max && !plan->fern && tile_fits) {
        // The cap says nothing about how orbit lengths are mixed inside a tile, and lock-step warps pay for the longest lane
        // (C3-like Julia dust: utilisation 0.38, 2.5x slower than the refill kernel).  The kernel reports its utilisation, so a
        // scene that did badly stays with the refill kernel for the next kTileBackoffFrames frames (harvest()).
        const uint64_t key = tile_key_of(cfg);
        if (c.tile_bad_key.load() == key && c.tile_bad_frames.load() > 0u) {
            c.tile_bad_frames.fetch_sub(1u);
        } else {
            v = FRB_KERNEL_TILE;
            plan->tile_key = key;
        }
    }
    plan->tile_form = form6_is_exact(cfg) ? 6 : 7;
    plan->tile_host = false;
    if (v == FRB_KERNEL_AUTO || v == FRB_KERNEL_DEFER6) v = deferred_is_safe(cfg) ? FRB_KERNEL_DEFER6 : FRB_KERNEL_DENSE6;
    if (v == FRB_KERNEL_DENSE6 && !form6_is_exact(cfg)) v = FRB_KERNEL_STREAM7;
    plan->variant = v;
    choose_deferred_shape(cfg, &plan->defer_b, &plan->defer_nb, &plan->defer_minb);
    p.defer_block = (uint32_t)plan->defer_b;
    // DEFER6 leaves provisional records; the shade kernel resolves them on the fly, the raw-result path (format 0) runs K1r
    p.records = (v == FRB_KERNEL_DEFER6 && format != 0) ? 1u : 0u;

    if (int rc = ensure(c.b().tab, c.b().tab_cap, (size_t)cfg.width + cfg.height + 2)) return rc;
    if (int rc = ensure(c.b().iters, c.b().iters_cap, (size_t)npix + 4)) return rc;
    if (int rc = ensure(c.b().zfinal, c.b().z_cap, (size_t)npix + 4)) return rc;
    p.tab_re = c.b().tab;
    p.tab_im = c.b().tab + cfg.width;
    p.iters = c.b().iters;
    p.zfinal = c.b().zfinal;
    p.queue = c.b().queue;
    p.stats = c.b().stats;
    return FRB_OK;
}

// CTAs of the deferred kernel for the chunks of `p` (persistent grid, never more than there are chunks for)
uint32_t deferred_blocks(const DeviceCtx &c, const Plan &plan, const RenderParams &p)
{
    const DeferShape sh = kDeferShapes[plan.defer_minb];
    const uint32_t warps_per_block = (uint32_t)sh.threads / 32;
    const uint32_t needed = (p.total_chunks + warps_per_block - 1) / warps_per_block;
    return std::max(1u, std::min((uint32_t)(c.sm_count * sh.per_sm), needed));
}

// The instantiation of the deferred kernel a plan selects (nullptr: none compiled for that combination).
const void *deferred_kernel_of(const Plan &plan, bool julia)
{
    const int key = plan.defer_b * 100 + plan.defer_nb * 10 + plan.defer_minb;
#define FRB_DEFER_CASE(B, NB, S)                                                                                                   \
    case (B) * 100 + (NB) * 10 + (S):                                                                                             \
        return julia ? (const void *)frb::frb_iterate_deferred_kernel<true, B, NB, kDeferShapes[S].per_sm, kDeferShapes[S].threads> \
                     : (const void *)frb::frb_iterate_deferred_kernel<false, B, NB, kDeferShapes[S].per_sm, kDeferShapes[S].threads>;
    switch (key) {
        // what choose_deferred_shape returns without tuning: caps < 256, < 2048, the rest; one wide CTA per SM
        FRB_DEFER_CASE(8, 1, 2) FRB_DEFER_CASE(8, 4, 2) FRB_DEFER_CASE(16, 4, 2)
#ifdef FRB_TUNING
        // the sweeps of profiles/*_tuning_sweeps.txt: other trip lengths with the default shape, other shapes with the default trips
        FRB_DEFER_CASE(8, 2, 2) FRB_DEFER_CASE(8, 8, 2) FRB_DEFER_CASE(16, 2, 2) FRB_DEFER_CASE(16, 8, 2) FRB_DEFER_CASE(32, 2, 2)
        FRB_DEFER_CASE(32, 4, 2)
        FRB_DEFER_CASE(8, 4, 0) FRB_DEFER_CASE(8, 4, 1) FRB_DEFER_CASE(8, 4, 3) FRB_DEFER_CASE(8, 4, 4) FRB_DEFER_CASE(8, 4, 5)
        FRB_DEFER_CASE(16, 4, 0) FRB_DEFER_CASE(16, 4, 1) FRB_DEFER_CASE(16, 4, 3) FRB_DEFER_CASE(16, 4, 4) FRB_DEFER_CASE(16, 4, 5)
#endif
    default:
        return nullptr;
    }
#undef FRB_DEFER_CASE
}

// CUDA loads a kernel's code on its first launch (lazy module loading), and that load waits for the kernels already running
// on the device.  The streaming finisher is launched FIRST and then sleeps until the frame's table kernel has run — so a table
// or iterate kernel that still has to be loaded at that point cannot start until the finisher gives up: 20 ms and the fallback
// path if it is the table kernel, a 4 s stall and an error if it is the iterate kernel (seen on the first streamed frame of
// each algorithm / block shape in a process).  Everything that is launched while a finisher may be resident is therefore
// loaded beforehand: cudaFuncGetAttributes forces the load.
int preload_kernel(DeviceCtx &c, const void *kfn)
{
    if (!kfn) return FRB_OK;
    for (const void *k : c.loaded_kernels)
        if (k == kfn) return FRB_OK;
    cudaFuncAttributes attr;
    CU(cudaFuncGetAttributes(&attr, kfn));
    c.loaded_kernels.push_back(kfn);
    return FRB_OK;
}

// Enqueues tables -> iterate -> shade (if format != 0) on `st`.  Caller holds c.mu and has set the device.
// Launches the iteration kernel chosen by plan.variant for the chunks [p.chunk_begin, p.chunk_begin + p.total_chunks).
int launch_iterate(DeviceCtx &c, const Plan &plan, const RenderParams &p, cudaStream_t st)
{
    if (plan.variant == FRB_KERNEL_SIMPLE) {
        dim3 grid(p.rows, (p.width + 255) / 256);
        frb::frb_iterate_simple_kernel<<<grid, 256, 0, st>>>(p);
        return FRB_OK;
    }
    if (plan.variant == FRB_KERNEL_TILE) {
        // iterates AND shades: the caller skips the shade kernel
        const size_t dyn = (size_t)std::min(44, std::max(0, tuning().tile_smem_kb)) << 10;
        const int tile_w = tuning().tile_w > 0 ? tuning().tile_w : (plan.tile_host ? kTileHostW : kTileDefaultW);
        const int key = tile_w * 100 + tuning().tile_wpc * 10 + (plan.tile_form == 6 ? 6 : 7);
#define FRB_TILE_CASE(TW, WPC)                                                                              \
    case (TW) * 100 + (WPC) * 10 + 6:                                                                       \
    case (TW) * 100 + (WPC) * 10 + 7: {                                                                     \
        dim3 grid((p.width + (TW) * (WPC) - 1) / ((TW) * (WPC)), (p.rows + 32 / (TW) - 1) / (32 / (TW)));   \
        if (grid.y > 65535u) return fail(FRB_ERR_TOO_LARGE, "tile kernel: %u rows exceed the grid's y extent", p.rows); \
        if (plan.tile_form == 6) frb::frb_iterate_tile_kernel<6, TW, WPC><<<grid, 32 * (WPC), dyn, st>>>(p); \
        else frb::frb_iterate_tile_kernel<7, TW, WPC><<<grid, 32 * (WPC), dyn, st>>>(p);                     \
        break;                                                                                              \
    }
        switch (key) {
            FRB_TILE_CASE(kTileDefaultW, kTileDefaultWpc) FRB_TILE_CASE(kTileHostW, kTileDefaultWpc)
#ifdef FRB_TUNING
            FRB_TILE_CASE(8, 1) FRB_TILE_CASE(8, 2) FRB_TILE_CASE(8, 8)
            FRB_TILE_CASE(16, 1) FRB_TILE_CASE(16, 2) FRB_TILE_CASE(16, 4)
            FRB_TILE_CASE(32, 1) FRB_TILE_CASE(32, 2)
#endif
        default:
            return fail(FRB_ERR_INVALID, "no tile kernel for %d-pixel-wide tiles, %d warps per CTA", tile_w, tuning().tile_wpc);
        }
#undef FRB_TILE_CASE
        return FRB_OK;
    }
    const uint32_t warps_per_block = frb::kIterateThreads / 32;
    if (plan.variant == FRB_KERNEL_DEFER6) {
        const uint32_t blocks = deferred_blocks(c, plan, p);
        const bool julia = (p.algo == FRB_ALGO_JULIA);
        const void *kfn = deferred_kernel_of(plan, julia);
        if (!kfn)
            return fail(FRB_ERR_INVALID, "no deferred kernel for block %d x %d, shape %d", plan.defer_b, plan.defer_nb, plan.defer_minb);
        // The queue of this launch (frb_deferred.cuh): rows [row0, row0 + nrows) of the shard; the last rows — about two refills
        // of every lane of the grid — go out in short chunks so that little unstarted work is left in any warp's hands when the
        // queue runs dry, and drained warps with few live lanes merge through the CTA's orbit pool.
        RenderParams q = p;
        {
            const uint32_t row0 = p.chunks_per_row ? p.chunk_begin / p.chunks_per_row : 0u;
            const uint32_t nrows = p.chunks_per_row ? p.total_chunks / p.chunks_per_row : 0u;
            const uint32_t threads = (uint32_t)kDeferShapes[plan.defer_minb].threads;
            // short chunks only pay where a pixel can be long (C2/8: 0.750 -> 0.711 ms with 4-pixel chunks and the pool; C3,
            // cap 1000: 0.909 -> 0.942 ms — the extra queue traffic costs more than the shorter tail gains)
            const int chunk2_default = p.iterations >= 2048u ? 4 : (int)p.chunk;
            const uint32_t chunk2 = (uint32_t)std::max(1, std::min(tuning().end_chunk > 0 ? tuning().end_chunk : chunk2_default, (int)p.chunk));
            const uint64_t end_px = 2ull * blocks * threads;
            uint32_t end_rows = (uint32_t)std::min<uint64_t>((end_px + p.width - 1) / std::max(p.width, 1u), nrows / 4u);
            if (chunk2 == p.chunk) end_rows = 0;
            q.q_row0 = row0;
            q.q_rows_big = nrows - end_rows;
            q.q_nbig = q.q_rows_big * p.chunks_per_row;
            q.q_chunk2 = chunk2;
            q.q_cpr2 = (p.width + chunk2 - 1) / chunk2;
            const uint64_t total = (uint64_t)q.q_nbig + (uint64_t)end_rows * q.q_cpr2;
            if (total > 0xfffffff0ull) {  // cannot happen below 2^32 pixels with >= 1-pixel chunks on a quarter of the rows; be safe
                q.q_rows_big = nrows;
                q.q_nbig = nrows * p.chunks_per_row;
                q.total_chunks = q.q_nbig;
            } else {
                q.total_chunks = (uint32_t)total;
            }
            q.q_pool_max = tuning().pool ? (uint32_t)std::min(31, std::max(0, tuning().pool_max)) : 0u;
        }
        void *args[] = {&q};
        CU(cudaLaunchKernel(kfn, dim3(blocks), dim3((unsigned)kDeferShapes[plan.defer_minb].threads), args, 0, st));
        // raw-result path: K1r over the rows this launch covered (with an image format the shade kernel resolves instead)
        RenderParams r = p;
        r.resolve_row0 = p.chunks_per_row ? p.chunk_begin / p.chunks_per_row : 0u;
        r.resolve_rows = p.chunks_per_row ? p.total_chunks / p.chunks_per_row : 0u;
        if (p.records == 0u && r.resolve_rows > 0) {
            dim3 grid(r.resolve_rows, (p.width + 255) / 256);
            frb::frb_resolve_kernel<<<grid, 256, 0, st>>>(r);
        }
        return FRB_OK;
    }
    const int kidx = (plan.variant == FRB_KERNEL_STREAM7) ? 1 : 0;
    uint32_t blocks = (uint32_t)(c.sm_count * c.iter_blocks_per_sm[kidx]);
    const uint32_t needed = (p.total_chunks + warps_per_block - 1) / warps_per_block;
    blocks = std::max(1u, std::min(blocks, needed));
    if (kidx == 0) frb::frb_iterate_stream_kernel<6><<<blocks, frb::kIterateThreads, 0, st>>>(p);
    else frb::frb_iterate_stream_kernel<7><<<blocks, frb::kIterateThreads, 0, st>>>(p);
    return FRB_OK;
}

// Shade (or, with format 0, only sum the iterations of) the pixels the pointers of `p` address: p.rows x p.width.
void launch_shade(DeviceCtx &c, const RenderParams &p, cudaStream_t st)
{
    const uint64_t npix = (uint64_t)p.rows * p.width;
    if (p.format != 0) {
        const uint64_t cta_px = (uint64_t)frb::kShadeThreads * frb::kShadePxPerThread;
        const uint64_t nblocks = (npix + cta_px - 1) / cta_px;
        const uint32_t grid = (uint32_t)std::min<uint64_t>(nblocks, (uint64_t)c.sm_count * 32);
        frb::frb_shade_kernel<<<grid, frb::kShadeThreads, 0, st>>>(p);
    } else {
        const uint32_t grid = (uint32_t)std::min<uint64_t>((npix + 255) / 256, (uint64_t)c.sm_count * 16);
        frb::frb_sum_iterations_kernel<<<grid, 256, 0, st>>>(p);
    }
}

Slot &take_slot(DeviceCtx &c, const Plan &plan)
{
    Slot &slot = c.slots[c.next_slot];
    c.next_slot = (c.next_slot + 1) % kSlots;
    slot.pixels = plan.npix;
    slot.launches = 0;
    slot.variant = plan.variant;
    slot.shaded = false;
    slot.stream_fallback = false;
    slot.tile_key = plan.variant == FRB_KERNEL_TILE ? plan.tile_key : 0;
    slot.host->iterations = 0;
    slot.host->lane_iterations = 0;
    slot.host->stream_status = 0;
    slot.host->stream_started = 0;
    return slot;
}

// Enqueues tables -> iterate -> shade (if format != 0) on `st`.  Caller holds c.mu and has set the device.
int enqueue(DeviceCtx &c, Plan &plan, uint8_t *dev_out, cudaStream_t st, Slot **slot_out)
{
    RenderParams &p = plan.p;
    p.out = dev_out;
    Slot &slot = take_slot(c, plan);
    *slot_out = &slot;

    // the raw buffers (tables, counts, final z, queue) are shared by all renders of this device: order streams
    if (c.b().have_guard && c.b().last_stream != st) CU(cudaStreamWaitEvent(st, c.b().ev_guard, 0));

    CU(cudaEventRecord(slot.ev_start, st));
    slot.have_k0 = false;
    if (plan.npix > 0) {
        const uint32_t n = std::max(p.width, p.height);
        frb::frb_coord_tables_kernel<<<(n + 255) / 256, 256, 0, st>>>(p);
        slot.launches++;
        if (!plan.fern) {
            CU(cudaEventRecord(slot.ev_k0, st));
            slot.have_k0 = true;
            if (int rc = launch_iterate(c, plan, p, st)) return rc;
            slot.launches += (plan.variant == FRB_KERNEL_DEFER6 && p.records == 0u) ? 2 : 1;  // raw DEFER6 = iterate + resolve
        } else {
            // get_recursive_pixel returns BLACK without iterating (calc/src/lib.rs:211); keep the raw outputs defined
            CU(cudaMemsetAsync(p.iters, 0, plan.npix * sizeof(uint32_t), st));
            CU(cudaMemsetAsync(p.zfinal, 0, plan.npix * sizeof(double2), st));
        }
    }
    CU(cudaEventRecord(slot.ev_iter, st));
    if (plan.npix > 0) {
        if (plan.variant != FRB_KERNEL_TILE) {  // the tile kernel has shaded already
            launch_shade(c, p, st);
            slot.launches++;
        }
        slot.shaded = (p.format != 0);
    }
    CU(cudaEventRecord(slot.ev_shade, st));
    if (plan.npix > 0) CU(cudaMemcpyAsync(slot.host, c.b().stats, sizeof(StatSlot), cudaMemcpyDeviceToHost, st));
    else memset(slot.host, 0, sizeof(StatSlot));  // an empty shard: nothing ran, nothing to report (not the slot's last frame)
    CU(cudaGetLastError());
    return FRB_OK;
}

// True when a blocking host-destination render should go through the streaming finisher (frb_stream.cuh): the deferred
// kernel, an image format and enough pixels for the second kernel to pay for itself.  Device destinations (frb_render_device_async) are not streamed: there is no copy to hide, the finisher's
// FP64 work competes with K1c's either way (measured: 4.84 ms fused shade vs 5.1-5.4 ms streamed on C2), and pipelined
// asynchronous frames would need one set of control words per frame in flight.
bool want_streamed(const Plan &plan)
{
    if (!tuning().stream || plan.fern) return false;
    return plan.variant == FRB_KERNEL_DEFER6 && plan.p.format != 0 && plan.npix >= (256u << 10) && plan.p.rows >= 64u;
}

// K2s on an auxiliary stream (launched first) | tables -> K1c on `st` -> join.  `dst` is the device alias of the caller's
// pinned host buffer: the finisher writes the image over PCIe while K1c iterates.  Blocking callers only (one frame in
// flight per device: the control words live in the device's statistics block).
// Slabs of the streaming finisher for the shard of `p`: ~256 Kpx each — small enough that the last slab (the only exposed one)
// is a few microseconds of shading and copying, large enough to give every finisher CTA a block or two.
struct StreamGeom {
    uint32_t slab_rows, nslabs, blocks_per_slab;
};
StreamGeom stream_geometry(const RenderParams &p)
{
    StreamGeom g;
    g.slab_rows = (uint32_t)std::max<uint64_t>(16u, ((256u << 10) / std::max(p.width, 1u) + 15u) & ~15u);
    if (tuning().stream_rows > 0) g.slab_rows = (uint32_t)std::max(16, (tuning().stream_rows + 15) & ~15);
    g.nslabs = (p.rows + g.slab_rows - 1) / g.slab_rows;
    const uint64_t cta_px = (uint64_t)frb::kShadeThreads * frb::kShadePxPerThread;
    g.blocks_per_slab = (uint32_t)(((uint64_t)g.slab_rows * p.width + cta_px - 1) / cta_px);
    return g;
}

int enqueue_streamed(DeviceCtx &c, Plan &plan, uint8_t *dst, cudaStream_t st, Slot **slot_out, uint32_t *slab_ready = nullptr)
{
    RenderParams &p = plan.p;
    p.out = dst;
    p.wm = c.b().wm;
    p.records = 1u;
    p.resolve_row0 = 0u;
    if (++c.stream_gen == 0u) ++c.stream_gen;
    p.stream_gen = c.stream_gen;
    Slot &slot = take_slot(c, plan);
    *slot_out = &slot;
    const uint32_t blocks = deferred_blocks(c, plan, p);
    const uint32_t nwarps = blocks * ((uint32_t)kDeferShapes[plan.defer_minb].threads / 32);
    if (nwarps > kMaxIterWarps) return fail(FRB_ERR_INVALID, "iterate grid of %u warps exceeds the watermark table", nwarps);

    frb::StreamCtl ctl;
    memset(&ctl, 0, sizeof ctl);
    ctl.wm = p.wm;
    // control words live in the 64-bit statistics slots: status + minimum in stats[2] (copied back with the statistics),
    // election + work counter in stats[10], generation in stats[11], the sticky failure flag in stats[12]
    ctl.words = reinterpret_cast<uint32_t *>(p.stats + 2);
    ctl.queue = reinterpret_cast<uint32_t *>(p.stats + 10);
    ctl.gen = reinterpret_cast<const uint32_t *>(p.stats + 11);
    ctl.started = reinterpret_cast<uint32_t *>(p.stats + 12);
    ctl.gen_value = p.stream_gen;
    ctl.gen_polls = (uint32_t)std::max(0, tuning().stream_gen_polls);  // ~20 ms (polls of ~1 us): K0 normally follows within microseconds
    ctl.nwarps = nwarps;
    // ~256 Kpx per slab: small enough that the last slab (the only exposed one) is a few microseconds of shading and copying,
    // large enough to give every finisher CTA a block or two
    const uint32_t slab_rows = stream_geometry(p).slab_rows;
    ctl.slab_rows = slab_rows;
    if (slab_ready != nullptr) {
        const StreamGeom geo = stream_geometry(p);
        ctl.slab_count = c.b().wm + kMaxIterWarps;  // the slab counters live behind the watermark table
        ctl.slab_ready = slab_ready;
        ctl.blocks_per_slab = geo.blocks_per_slab;
    }
    {
        // bound on one wait, in polls of >= 2 us: a few seconds plus a generous allowance for the slab's own work, so that it
        // is only ever reached if K1c stopped making progress
        const uint64_t work = (uint64_t)slab_rows * p.width * (uint64_t)std::max(p.iterations, 1u);
        ctl.max_polls = (uint32_t)std::min<uint64_t>((2u << 20) + (work >> 22), 0x7fffffffu);
    }
    const int per_sm = std::min(8, std::max(1, tuning().stream_ctas));
    // The finisher goes first and depends on nothing (it sleeps on the generation word until K0 has run), so that its CTAs are
    // resident before — older than — K1c's.  Frames alternate between the two auxiliary streams so that the next frame's
    // finisher can take its seat while this frame is still running.
    if (int rc = preload_kernel(c, (const void *)frb::frb_coord_tables_kernel)) return rc;
    if (int rc = preload_kernel(c, deferred_kernel_of(plan, p.algo == FRB_ALGO_JULIA))) return rc;
    if (int rc = preload_kernel(c, (const void *)frb::frb_finish_stream_kernel)) return rc;
    // (the watermark reset goes ahead of the finisher too: whatever a memset needs is then in place before anything sleeps on
    // the device; the finisher does not look at the table before K0, which follows the reset on `st`, has published the frame)
    if (c.b().have_guard && c.b().last_stream != st) CU(cudaStreamWaitEvent(st, c.b().ev_guard, 0));
    CU(cudaEventRecord(slot.ev_start, st));
    CU(cudaMemsetAsync(p.wm, 0, nwarps * sizeof(uint32_t), st));
    if (slab_ready != nullptr) CU(cudaMemsetAsync(ctl.slab_count, 0, stream_geometry(p).nslabs * sizeof(uint32_t), st));
    cudaStream_t fs = c.aux[p.stream_gen & 1u];
    frb::frb_finish_stream_kernel<<<(unsigned)(c.sm_count * per_sm), frb::kShadeThreads, 0, fs>>>(p, ctl);
    CU(cudaEventRecord(c.ev_join[p.stream_gen & 1u], fs));

    const uint32_t n = std::max(p.width, p.height);
    frb::frb_coord_tables_kernel<<<(n + 255) / 256, 256, 0, st>>>(p);
    CU(cudaEventRecord(slot.ev_k0, st));
    slot.have_k0 = true;
    if (int rc = launch_iterate(c, plan, p, st)) return rc;
    CU(cudaEventRecord(slot.ev_iter, st));
    CU(cudaStreamWaitEvent(st, c.ev_join[p.stream_gen & 1u], 0));
    slot.launches += 3;
    slot.shaded = true;
    CU(cudaEventRecord(slot.ev_shade, st));
    CU(cudaMemcpyAsync(slot.host, c.b().stats, sizeof(StatSlot), cudaMemcpyDeviceToHost, st));
    CU(cudaGetLastError());
    return FRB_OK;
}

// Device alias of a host pointer the GPU can write (pinned / registered memory under UVA), or nullptr.
uint8_t *device_alias_of_host(const void *host)
{
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, host) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    if (attr.type != cudaMemoryTypeHost || attr.devicePointer == nullptr) return nullptr;
    return static_cast<uint8_t *>(attr.devicePointer);
}

// Host-output path for large frames: the shard is cut into row slabs, each iterated + shaded + copied out on one of two
// auxiliary streams.  The iterate kernels are persistent full-GPU grids, so slab s+1's CTAs start as slab s's CTAs
// drain (its tail is filled) and slab s's device->host copy runs under slab s+1's iterations.  Nothing waits on
// anything but stream order and events.  `host_out` should be pinned for the copies to be asynchronous.
// Copies the shard's local rows [r0, r1) (r0 a multiple of the band height) from `src` to the host.  band_pitch == 0: the
// destination is the shard's own compact image.  Otherwise the destination is a whole frame in which this shard's band b starts
// at host_out + b * band_pitch (interleaved bands of band_rows rows: frb_render with device = -1, frb_render_bands): one strided
// copy for the full bands, one more for a partial last band.
int copy_rows_out(uint8_t *host_out, size_t band_pitch, uint32_t band_rows, const uint8_t *src, uint32_t r0, uint32_t r1, size_t row_bytes,
                  cudaStream_t cs)
{
    if (r1 <= r0) return FRB_OK;
    if (band_pitch == 0) {
        CU(cudaMemcpyAsync(host_out + (size_t)r0 * row_bytes, src, (size_t)(r1 - r0) * row_bytes, cudaMemcpyDeviceToHost, cs));
        return FRB_OK;
    }
    const size_t band_bytes = row_bytes * band_rows;
    const uint32_t b0 = r0 / band_rows;
    const uint32_t full = (r1 - r0) / band_rows;
    uint8_t *dst0 = host_out + (size_t)b0 * band_pitch;
    if (full > 0) CU(cudaMemcpy2DAsync(dst0, band_pitch, src, band_bytes, band_bytes, full, cudaMemcpyDeviceToHost, cs));
    const uint32_t rest = (r1 - r0) - full * band_rows;
    if (rest > 0)
        CU(cudaMemcpyAsync(dst0 + (size_t)full * band_pitch, src + (size_t)full * band_bytes, (size_t)rest * row_bytes, cudaMemcpyDeviceToHost, cs));
    return FRB_OK;
}

// Rows per slab of the sliced host path (equal slabs, multiples of 16 rows so that every slab starts 16-byte aligned).
uint32_t sliced_rows_per(const RenderParams &p, int nslabs, size_t band_pitch)
{
    uint32_t rows_per = (p.rows + (uint32_t)nslabs - 1) / (uint32_t)nslabs;
    rows_per = std::max(16u, (rows_per + 15u) & ~15u);
    if (band_pitch != 0) rows_per = (rows_per + p.band_rows - 1) / p.band_rows * p.band_rows;  // slabs of whole bands
    return rows_per;
}

void CUDART_CB raise_flag(void *flag) { *static_cast<volatile uint32_t *>(flag) = 1u; }

// `ready` != nullptr: host_out is the library's pinned staging buffer (compact image) and ready[s] is raised — by a host function
// in stream order — once slab s has arrived there (pageable destinations: a CopyJob moves it on from there).
int enqueue_sliced(DeviceCtx &c, Plan &plan, uint8_t *host_out, int nslabs, Slot **slot_out, size_t band_pitch = 0,
                   uint32_t *ready = nullptr)
{
    RenderParams &p = plan.p;
    p.out = c.b().out;
    Slot &slot = take_slot(c, plan);
    *slot_out = &slot;
    cudaStream_t st = c.stream;
    if (c.b().have_guard && c.b().last_stream != st) CU(cudaStreamWaitEvent(st, c.b().ev_guard, 0));
    CU(cudaEventRecord(slot.ev_start, st));
    slot.have_k0 = false;
    const uint32_t n = std::max(p.width, p.height);
    frb::frb_coord_tables_kernel<<<(n + 255) / 256, 256, 0, st>>>(p);
    slot.launches++;
    CU(cudaEventRecord(c.ev_fork, st));
    // Equal slabs (multiples of 16 rows keep 16-byte alignment), alternating between the two auxiliary streams so that their
    // kernels overlap freely.  Measured and rejected (profiles/r1_tuning_sweeps.txt): slabs that shrink towards the end of the
    // frame (C2 5.267 vs 5.244 ms, C3 1.711 vs 1.632, C4 210.5 vs 210.8); kernels on one stream with the copies on the other
    // (C2 5.58-5.78 vs 5.555 ms).
    int s = 0;
    for (uint32_t r0 = 0; r0 < p.rows; ++s) {
        const uint32_t rows_per = sliced_rows_per(p, nslabs, band_pitch);
        cudaStream_t ks = c.aux[s & 1];
        cudaStream_t cs = ks;
        if (s < 2) CU(cudaStreamWaitEvent(c.aux[s & 1], c.ev_fork, 0));
        const uint32_t r1 = (s >= nslabs - 1) ? p.rows : std::min(p.rows, r0 + rows_per);
        RenderParams it = p;  // iterate: global pixel indexing, its own chunk range and queue counter
        it.chunk_begin = r0 * p.chunks_per_row;
        it.total_chunks = (r1 - r0) * p.chunks_per_row;
        it.queue = c.b().queue + (s % 16);
        RenderParams sh = p;  // shade: the slab's pixels through offset pointers
        const size_t pix0 = (size_t)r0 * p.width;
        sh.rows = r1 - r0;
        sh.resolve_row0 = r0;
        sh.iters = p.iters + pix0;
        sh.zfinal = p.zfinal + pix0;
        frb::offset_out(sh, p, pix0);
        if (plan.variant == FRB_KERNEL_TILE) {
            // one kernel iterates and shades; it addresses the slab the way the shade kernel does
            if (int rc = launch_iterate(c, plan, sh, ks)) return rc;
            slot.launches += 1;
        } else {
            if (int rc = launch_iterate(c, plan, it, ks)) return rc;
            launch_shade(c, sh, ks);
            slot.launches += 2;
        }
        if (int rc = copy_rows_out(host_out, band_pitch, p.band_rows, sh.out, r0, r1, (size_t)p.width * p.format, cs)) return rc;
        if (ready != nullptr) CU(cudaLaunchHostFunc(cs, raise_flag, ready + s));
        r0 = r1;
    }
    for (int i = 0; i < 2; ++i) {
        CU(cudaEventRecord(c.ev_join[i], c.aux[i]));
        CU(cudaStreamWaitEvent(st, c.ev_join[i], 0));
    }
    slot.shaded = true;
    CU(cudaEventRecord(slot.ev_iter, st));
    CU(cudaEventRecord(slot.ev_shade, st));
    CU(cudaMemcpyAsync(slot.host, c.b().stats, sizeof(StatSlot), cudaMemcpyDeviceToHost, st));
    CU(cudaGetLastError());
    return FRB_OK;
}

int finish_enqueue(DeviceCtx &c, Slot &slot, cudaStream_t st)
{
    CU(cudaEventRecord(slot.ev_end, st));
    CU(cudaEventRecord(c.b().ev_guard, st));
    c.b().have_guard = true;
    c.b().last_stream = st;
    c.b().last_use = ++c.use_tick;
    return FRB_OK;
}

int harvest(DeviceCtx &c, Slot &slot, FrbStats *out, double total_ms)
{
    CU(cudaEventSynchronize(slot.ev_end));
    if ((uint32_t)slot.host->stream_status != 0u && !slot.stream_fallback)
        return fail(FRB_ERR_CUDA, "the streaming finisher gave up waiting for the iterate kernel (frb_stream.cuh: stream_wait)");
    FrbStats s = {};
    s.pixels = slot.pixels;
    s.iterations = slot.host->iterations;
    s.lane_iterations = slot.host->lane_iterations;
    if (slot.variant == FRB_KERNEL_TILE)
        for (uint32_t k = 0; k < frb::kTileStatsPairs; ++k) {
            s.iterations += slot.host->tile_part[2 * k];
            s.lane_iterations += slot.host->tile_part[2 * k + 1];
        }
    if (slot.tile_key != 0 && slot.pixels >= (1u << 16) && s.lane_iterations > 0 &&
        (double)s.iterations < kTileMinUtilisation * (double)s.lane_iterations) {
        c.tile_bad_key.store(slot.tile_key);
        c.tile_bad_frames.store(kTileBackoffFrames);
    }
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, slot.have_k0 ? slot.ev_k0 : slot.ev_start, slot.ev_iter));  // the iterate kernel alone where K0 has its own mark
    s.iterate_ms = ms;
    CU(cudaEventElapsedTime(&ms, slot.ev_iter, slot.ev_shade));
    s.shade_ms = ms;
    CU(cudaEventElapsedTime(&ms, slot.ev_start, slot.ev_shade));
    s.device_ms = ms;
    s.total_ms = (float)total_ms;
    s.kernel_launches = slot.launches;
    s.devices = 1;
    s.kernel_variant = slot.variant;
    *out = s;
    return FRB_OK;
}

double now_ms()
{
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

#include "frb_staging.h"  // StagePool / CopyJob / CopyPool: pageable destinations through pinned staging and host copy threads

// One blocking render of one shard on one device into a host buffer (RGB/RGBA) and/or raw host buffers.
// framed = false: `host_out` receives the shard's own compact image (its rows in increasing y).
// framed = true:  `host_out` is the WHOLE frame and the shard's bands go to their place in it (band b of the shard starts at
//                 ((b * count + index) * band_rows) rows) — how several devices, or several processes, assemble one image.
// Error paths: work already launched may still write into a staging buffer (image or flags) or the caller's memory.  Wait for
// the device's three streams — ignoring what they report, the caller already has its error — before that memory changes hands.
void drain_streams(DeviceCtx &c)
{
    cudaStreamSynchronize(c.stream);
    cudaStreamSynchronize(c.aux[0]);
    cudaStreamSynchronize(c.aux[1]);
    cudaGetLastError();
}

int render_blocking(const FrbConfig &cfg, const FrbShard &sh, int dev, uint32_t format, uint8_t *host_out, size_t out_len,
                    uint32_t *host_iters, double *host_z, bool framed = false)
{
    const double t0 = now_ms();
    DeviceCtx *c = get_ctx(dev);
    if (!c) return tl_code;
    c->waiting.fetch_add(1, std::memory_order_relaxed);
    std::lock_guard<std::mutex> lk(c->mu);
    c->waiting.fetch_sub(1, std::memory_order_relaxed);
    CU(cudaSetDevice(dev));
    Plan plan;
    if (int rc = make_plan(*c, cfg, sh, format, &plan)) return rc;
    const size_t row_bytes = (size_t)cfg.width * format;
    const size_t band_bytes = row_bytes * plan.p.band_rows;
    if (sh.count <= 1) framed = false;  // the whole frame is its own compact image
    if (format != 0) {
        const size_t need = framed ? row_bytes * cfg.height : plan.out_bytes;
        if (out_len < need) return fail(FRB_ERR_INVALID, "output buffer too small: %zu < %zu bytes", out_len, need);
        if (plan.out_bytes && !host_out) return fail(FRB_ERR_INVALID, "null output buffer");
        if (int rc = ensure(c->b().out, c->b().out_cap, plan.out_bytes + 64)) return rc;
    }
    // where band 0 of the shard starts, and the distance between two of its bands (0 = compact)
    uint8_t *const dst0 = framed ? host_out + (size_t)sh.index * band_bytes : host_out;
    const size_t band_pitch = framed ? (size_t)sh.count * band_bytes : 0;
    Slot *slot = nullptr;
    // Large RGB/RGBA frames go out in row slabs so that the device->host copy hides under the iterations.
    // The tile kernel's frames are short (tens of microseconds a megapixel), so the copy dominates from ~0.5 Mpx on.
    const bool tile = plan.variant == FRB_KERNEL_TILE;
    const bool image_only = format != 0 && !host_iters && !host_z && !plan.fern;
    const bool sliced = image_only && plan.variant != FRB_KERNEL_SIMPLE && plan.npix >= (tile ? (1u << 19) : (4u << 20)) &&
                        plan.p.rows >= 256u && !tuning().no_slab;
    uint8_t *alias = nullptr;
    if (image_only && want_streamed(plan) && (!framed || ((band_bytes & 15u) == 0 && row_bytes < (1ull << 32)))) {
        alias = device_alias_of_host(host_out);
        if (alias && (reinterpret_cast<uintptr_t>(alias) & 15u) != 0) alias = nullptr;  // the shade stores are 16-byte vectors
    }
    // Tile kernel + pinned compact destination: the kernel's word stores go straight into the caller's memory over PCIe (one
    // launch, no device image, no copy) instead of four slab launches each followed by a copy.
    uint8_t *tile_alias = nullptr;
    // (C1, 2 Mpx: 0.229 -> 0.210 ms; at 8 Mpx the narrow stores saturate the link: 0.574 -> 0.688 ms, so slabs stay beyond 3 Mpx)
    if (tile && image_only && !framed && tuning().tile_direct && plan.npix <= (3u << 20)) {
        tile_alias = device_alias_of_host(host_out);
        if (tile_alias && (reinterpret_cast<uintptr_t>(tile_alias) & 3u) != 0) tile_alias = nullptr;
    }
    StageBuf *tile_stage = nullptr;
    if (tile && image_only && !tile_alias && tuning().tile_direct && tuning().stage && plan.npix >= (1u << 19) && plan.npix <= (3u << 20) &&
        device_alias_of_host(host_out) == nullptr)
        tile_stage = g_stage.acquire(plan.out_bytes);
    StageBuf *sb = nullptr;
    if (!alias && image_only && want_streamed(plan) && tuning().stage && plan.out_bytes <= kMaxStageBytes &&
        stream_geometry(plan.p).nslabs <= kMaxStageSlabs)
        sb = g_stage.acquire(plan.out_bytes);
    if (sb) {
        // pageable destination: the finisher writes the shard's compact image into pinned staging, host threads copy slab after
        // slab into the caller's memory while the iterate kernel runs (see StagePool / CopyJob above)
        const StreamGeom geo = stream_geometry(plan.p);
        memset(sb->done, 0, (size_t)geo.nslabs * sizeof(uint32_t));
        auto job = std::make_shared<CopyJob>();
        job->stage = sb->p;
        job->dst = host_out;
        job->row_bytes = row_bytes;
        job->rows = plan.p.rows;
        job->slab_rows = geo.slab_rows;
        job->nslabs = geo.nslabs;
        job->blocks_per_slab = geo.blocks_per_slab;
        job->done = sb->done;
        job->framed = framed;
        job->band_rows = plan.p.band_rows;
        job->shard_index = sh.index;
        job->shard_count = sh.count;
        const int helpers = g_copy.helpers();
        job->running.store(helpers + 1);
        advise_huge_pages(framed ? host_out : dst0, framed ? row_bytes * cfg.height : plan.out_bytes);
        job->plan_populate();
        int rc = enqueue_streamed(*c, plan, sb->p, c->stream, &slot, sb->done);
        if (!rc) rc = finish_enqueue(*c, *slot, c->stream);
        if (rc) {
            drain_streams(*c);  // the finisher / the kernels launched before the failing call may still be writing
            g_stage.release(sb);
            return rc;
        }
        g_copy.post(job, helpers);
        const cudaError_t se = cudaStreamSynchronize(c->stream);
        const bool bad = se != cudaSuccess || (uint32_t)slot->host->stream_started == 0u || (uint32_t)slot->host->stream_status != 0u;
        if (bad) job->abort.store(true);
        job->run();  // the rendering thread helps with what is left (returns at once when aborted)
        job->wait_idle();
        g_stage.release(sb);
        if (se != cudaSuccess) return fail(FRB_ERR_CUDA, "cudaStreamSynchronize failed: %s", cudaGetErrorString(se));
        if (bad) {
            // the finisher never saw the frame start, or gave up (same recovery as for a pinned destination below)
            RenderParams sp = plan.p;
            sp.out = c->b().out;
            CU(cudaMemsetAsync(sp.stats, 0, sizeof(unsigned long long), c->stream));
            launch_shade(*c, sp, c->stream);
            slot->launches++;
            slot->stream_fallback = true;
            CU(cudaEventRecord(slot->ev_shade, c->stream));
            if (int rc2 = copy_rows_out(dst0, band_pitch, plan.p.band_rows, c->b().out, 0u, plan.p.rows, row_bytes, c->stream)) return rc2;
            CU(cudaMemcpyAsync(slot->host, c->b().stats, sizeof(StatSlot), cudaMemcpyDeviceToHost, c->stream));
        }
    } else if (alias) {
        // pinned destination: the finisher writes the image into the caller's buffer while the iterate kernel runs; with a
        // framed destination through strided output addressing (RenderParams::out_band_q) — no device image, no copy
        if (framed) {
            plan.p.out_band_q = (uint32_t)(band_bytes >> 4);
            plan.p.out_band_gap = (unsigned long long)(sh.count - 1) * band_bytes;
            plan.p.out_off = 0;
            alias += (size_t)sh.index * band_bytes;
        }
        if (int rc = enqueue_streamed(*c, plan, alias, c->stream, &slot)) return rc;
        if (int rc = finish_enqueue(*c, *slot, c->stream)) return rc;
        CU(cudaStreamSynchronize(c->stream));
        if ((uint32_t)slot->host->stream_started == 0u || (uint32_t)slot->host->stream_status != 0u) {
            // No finisher CTA saw the frame start (a profiler serialising kernels, or > 20 ms of other work ahead of this
            // frame), or one gave up waiting part-way (a co-tenant holding the iterate kernel back): the provisional records
            // are all in device memory, so shade them now with the ordinary kernel and copy the image out.
            RenderParams sp = plan.p;
            sp.out = c->b().out;
            sp.out_band_q = 0u;
            sp.out_off = 0;
            CU(cudaMemsetAsync(sp.stats, 0, sizeof(unsigned long long), c->stream));  // the iteration total is summed again
            launch_shade(*c, sp, c->stream);
            slot->launches++;
            slot->stream_fallback = true;
            CU(cudaEventRecord(slot->ev_shade, c->stream));
            if (int rc = copy_rows_out(dst0, band_pitch, plan.p.band_rows, c->b().out, 0u, plan.p.rows, row_bytes, c->stream)) return rc;
            CU(cudaMemcpyAsync(slot->host, c->b().stats, sizeof(StatSlot), cudaMemcpyDeviceToHost, c->stream));
        }
    } else if (tile_alias) {
        plan.tile_host = true;
        if (int rc = enqueue(*c, plan, tile_alias, c->stream, &slot)) return rc;
    } else if (tile_stage) {
        // Tile kernel + small pageable destination: ONE launch whose stores go straight into pinned staging over PCIe (as for
        // a pinned destination) while the copy threads give the destination its pages; then the threads copy.  Eight slab
        // launches, copies and host callbacks (the sliced path below) cost more than the frame: C1 fresh 1.07 -> 0.4x ms.
        const uint32_t rows_per = std::max(1u, (plan.p.rows + 7u) / 8u);
        const uint32_t nsl = (plan.p.rows + rows_per - 1) / rows_per;
        memset(tile_stage->done, 0, (size_t)nsl * sizeof(uint32_t));
        auto job = std::make_shared<CopyJob>();
        job->stage = tile_stage->p;
        job->dst = host_out;
        job->row_bytes = row_bytes;
        job->rows = plan.p.rows;
        job->slab_rows = rows_per;
        job->nslabs = nsl;
        job->done = tile_stage->done;
        job->framed = framed;
        job->band_rows = plan.p.band_rows;
        job->shard_index = sh.index;
        job->shard_count = sh.count;
        const int helpers = g_copy.helpers();
        job->running.store(helpers + 1);
        advise_huge_pages(framed ? host_out : dst0, framed ? row_bytes * cfg.height : plan.out_bytes);
        job->plan_populate();
        plan.tile_host = true;
        int rc = enqueue(*c, plan, tile_stage->p, c->stream, &slot);
        if (!rc) rc = finish_enqueue(*c, *slot, c->stream);
        if (rc) {
            drain_streams(*c);  // the finisher / the kernels launched before the failing call may still be writing
            g_stage.release(tile_stage);
            return rc;
        }
        g_copy.post(job, helpers);
        const cudaError_t se = cudaStreamSynchronize(c->stream);
        if (se != cudaSuccess) job->abort.store(true);
        std::atomic_thread_fence(std::memory_order_release);
        for (uint32_t k = 0; k < nsl; ++k) tile_stage->done[k] = 1u;  // the whole image is in staging
        job->run();
        job->wait_idle();
        g_stage.release(tile_stage);
        if (se != cudaSuccess) return fail(FRB_ERR_CUDA, "cudaStreamSynchronize failed: %s", cudaGetErrorString(se));
    } else if (sliced) {
        // every slab is one more persistent launch (~0.1 ms of ramp and tail) but exposes a smaller last copy: 4 slabs up to
        // 32 Mpx, 8 up to 64 Mpx, 16 beyond (C2 5.24 ms with 4; C4 213.5 / 211.4 / 210.5 ms with 4 / 8 / 16)
        int nslabs = plan.npix >= (64u << 20) ? 16 : (plan.npix >= (32u << 20) ? 8 : 4);
        if (tuning().slabs > 0) nslabs = std::min(16, tuning().slabs);
        // A pageable destination gets its slabs through pinned staging and the copy threads as well (the tile kernel's frames,
        // the dense kernels'): device -> staging at DMA speed, staging -> destination by several threads, first touch included.
        StageBuf *sb2 = nullptr;
        if (tuning().stage && plan.out_bytes <= kMaxStageBytes && device_alias_of_host(host_out) == nullptr) sb2 = g_stage.acquire(plan.out_bytes);
        if (sb2) {
            if (tile && nslabs < 8) nslabs = 8;  // short frames: smaller slabs start the host copies earlier
            const uint32_t rows_per = sliced_rows_per(plan.p, nslabs, 0);
            const uint32_t nsl = (plan.p.rows + rows_per - 1) / rows_per;
            memset(sb2->done, 0, (size_t)nsl * sizeof(uint32_t));
            auto job = std::make_shared<CopyJob>();
            job->stage = sb2->p;
            job->dst = host_out;
            job->row_bytes = row_bytes;
            job->rows = plan.p.rows;
            job->slab_rows = rows_per;
            job->nslabs = nsl;
            job->done = sb2->done;
            job->framed = framed;
            job->band_rows = plan.p.band_rows;
            job->shard_index = sh.index;
            job->shard_count = sh.count;
            const int helpers = g_copy.helpers();
            job->running.store(helpers + 1);
            advise_huge_pages(framed ? host_out : dst0, framed ? row_bytes * cfg.height : plan.out_bytes);
            job->plan_populate();
            int rc = enqueue_sliced(*c, plan, sb2->p, nslabs, &slot, 0, sb2->done);
            if (rc) {
                drain_streams(*c);  // slab copies and host functions queued before the failing call
                g_stage.release(sb2);
                return rc;
            }
            g_copy.post(job, helpers);
            job->run(c->stream);  // the flags are raised in stream order: the rendering thread copies along, and watches the stream
            job->wait_idle();
            if (job->abort.load()) {
                const cudaError_t we = job->watch_error;
                drain_streams(*c);
                g_stage.release(sb2);
                return fail(FRB_ERR_CUDA, "sliced render abandoned: %s", cudaGetErrorString(we));
            }
            g_stage.release(sb2);
        } else if (int rc = enqueue_sliced(*c, plan, dst0, nslabs, &slot, band_pitch)) {
            return rc;
        }
    } else {
        if (int rc = enqueue(*c, plan, c->b().out, c->stream, &slot)) return rc;
        if (plan.npix > 0) {
            if (format != 0)
                if (int rc = copy_rows_out(dst0, band_pitch, plan.p.band_rows, c->b().out, 0u, plan.p.rows, row_bytes, c->stream)) return rc;
            if (host_iters) CU(cudaMemcpyAsync(host_iters, c->b().iters, plan.npix * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
            if (host_z) CU(cudaMemcpyAsync(host_z, c->b().zfinal, plan.npix * sizeof(double2), cudaMemcpyDeviceToHost, c->stream));
        }
    }
    if (int rc = finish_enqueue(*c, *slot, c->stream)) return rc;
    CU(cudaStreamSynchronize(c->stream));
    return harvest(*c, *slot, &tl_stats, now_ms() - t0);
}

struct AsyncRecord {
    int dev = -1;
    Slot *slot = nullptr;
    double t0 = 0.0;
};
thread_local AsyncRecord tl_async_ring[kSlots];
thread_local uint64_t tl_async_count = 0;

#include "frb_workers.h"  // DeviceWorker, get_worker, stop_workers_locked

std::vector<int> active_devices()
{
    std::vector<int> v, list;
    int n;
    uint32_t mask;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        n = device_count_locked();
        mask = g_device_mask;
        list = g_device_list;
    }
    if (!list.empty()) return list;  // frb_init_list
    for (int d = 0; d < n; ++d)
        if (mask == 0 || (mask >> d) & 1u) v.push_back(d);
    return v;
}

}  // namespace

// frb_last_error() for the other translation units of the library (frb_shared.cu)
int frb_internal_fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    return fail(code, "%s", buf);
}

// ============================================================ C ABI ============================================================

extern "C" {

uint32_t frb_abi_version(void) { return FRB_ABI_VERSION; }

void *frb_host_alloc(size_t bytes)
{
    if (frb_device_count() == 0) {
        fail(FRB_ERR_NO_DEVICE, "no usable CUDA device: pinned memory needs the driver");
        return nullptr;
    }
    void *p = nullptr;
    const cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) {
        cudaGetLastError();
        fail(FRB_ERR_CUDA, "cudaHostAlloc(%zu bytes): %s", bytes, cudaGetErrorString(e));
        return nullptr;
    }
    return p;
}

void frb_host_free(void *p)
{
    if (p && cudaFreeHost(p) != cudaSuccess) cudaGetLastError();
}

int frb_host_register(void *p, size_t bytes)
{
    if (!p || bytes == 0) return fail(FRB_ERR_INVALID, "null or empty buffer");
    if (frb_device_count() == 0) return fail(FRB_ERR_NO_DEVICE, "no usable CUDA device: pinned memory needs the driver");
    CU(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
    return FRB_OK;
}

int frb_host_unregister(void *p)
{
    if (!p) return fail(FRB_ERR_INVALID, "null buffer");
    CU(cudaHostUnregister(p));
    return FRB_OK;
}

const char *frb_last_error(void) { return tl_error.c_str(); }

FrbRGB frb_rgb_new(uint8_t r, uint8_t b, uint8_t g)
{
    FrbRGB c;
    c.r = r;
    c.g = g;
    c.b = b;
    return c;
}

void frb_config_default(FrbConfig *cfg, uint32_t algo)
{
    if (!cfg) return;
    memset(cfg, 0, sizeof *cfg);
    const bool fern = (algo == FRB_ALGO_BARNSLEY_FERN);
    cfg->algo = algo;
    cfg->width = 2000;
    cfg->height = 1000;
    cfg->iterations = fern ? 10000000u : 50u;
    cfg->limit = 65536.0;
    cfg->stable_limit = 2.0;
    cfg->scale.re = 0.4;
    cfg->scale.im = 0.4;
    cfg->exposure = 2.0;
    cfg->inside = 1;
    cfg->smooth = 1;
    cfg->primary_color = fern ? frb_rgb_new(4, 100, 3) : frb_rgb_new(40, 40, 255);
    cfg->secondary_color = fern ? frb_rgb_new(240, 240, 240) : frb_rgb_new(240, 170, 0);
    cfg->color_weight = 0.01;
}

uint32_t frb_shard_rows(uint32_t height, const FrbShard *shard)
{
    const FrbShard s = shard ? *shard : whole_frame();
    return shard_rows(height, s);
}

uint32_t frb_shard_row_to_y(uint32_t height, const FrbShard *shard, uint32_t local_row)
{
    const FrbShard s = shard ? *shard : whole_frame();
    if (!shard_valid(s)) return 0;
    return shard_row_to_y(height, s, local_row);
}

int frb_device_count(void)
{
    std::lock_guard<std::mutex> lk(g_mu);
    return device_count_locked();
}

// Creates the context of device d and loads the code of the kernels an ordinary frame launches (CUDA would otherwise do it at
// their first launch, tens of milliseconds into the first render): tables, shade, finisher, tile, and the default shapes of the
// deferred kernel.
static int warm_device(int d)
{
    DeviceCtx *c = get_ctx(d);
    if (!c) return tl_code;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(d));
    const void *common[] = {(const void *)frb::frb_coord_tables_kernel, (const void *)frb::frb_shade_kernel,
                            (const void *)frb::frb_finish_stream_kernel, (const void *)frb::frb_iterate_tile_kernel<6, kTileDefaultW, kTileDefaultWpc>,
                            (const void *)frb::frb_iterate_tile_kernel<6, kTileHostW, kTileDefaultWpc>,
                            (const void *)frb::frb_resolve_kernel, (const void *)frb::frb_sum_iterations_kernel};
    for (const void *k : common)
        if (int rc = preload_kernel(*c, k)) return rc;
    for (uint32_t it : {64u, 1000u, 4000u})
        for (int julia = 0; julia < 2; ++julia) {
            FrbConfig probe = {};
            probe.iterations = it;
            Plan pl = {};
            choose_deferred_shape(probe, &pl.defer_b, &pl.defer_nb, &pl.defer_minb);
            if (int rc = preload_kernel(*c, deferred_kernel_of(pl, julia != 0))) return rc;
        }
    return FRB_OK;
}

int frb_init(uint32_t device_mask)
{
    int n = frb_device_count();
    if (n == 0) return fail(FRB_ERR_NO_DEVICE, "no usable CUDA device; this library has no CPU path");
    bool any = false;
    for (int d = 0; d < n; ++d) {
        if (device_mask != 0 && !((device_mask >> d) & 1u)) continue;
        if (int rc = warm_device(d)) return rc;
        any = true;
    }
    if (!any) return fail(FRB_ERR_INVALID, "device mask 0x%x selects none of the %d devices", device_mask, n);
    std::lock_guard<std::mutex> lk(g_mu);
    g_device_mask = device_mask;
    g_device_list.clear();
    return FRB_OK;
}

int frb_init_list(const int32_t *devices, uint32_t n)
{
    const int have = frb_device_count();
    if (have == 0) return fail(FRB_ERR_NO_DEVICE, "no usable CUDA device; this library has no CPU path");
    if (!devices || n == 0 || n > (uint32_t)kMaxDevices) return fail(FRB_ERR_INVALID, "device list of %u entries (1..%d)", n, kMaxDevices);
    for (uint32_t i = 0; i < n; ++i) {
        if (devices[i] < 0 || devices[i] >= have) return fail(FRB_ERR_INVALID, "device %d out of range [0, %d)", devices[i], have);
        if (int rc = warm_device(devices[i])) return rc;
    }
    std::lock_guard<std::mutex> lk(g_mu);
    g_device_list.assign(devices, devices + n);
    g_device_mask = 0;
    return FRB_OK;
}

int frb_set_tuning(const char *name, int64_t value)
{
    if (!name) return fail(FRB_ERR_INVALID, "null tuning name");
#ifdef FRB_TUNING
    using Field = int Tuning::*;
    struct Entry {
        const char *name;
        Field field;
    };
    static const Entry table[] = {
        {"defer_b", &Tuning::defer_b},       {"defer_nb", &Tuning::defer_nb},
        {"defer_shape", &Tuning::defer_shape}, {"chunk", &Tuning::chunk},
        {"fast_shade", &Tuning::fast_shade}, {"tile_max", &Tuning::tile_max},
        {"stream", &Tuning::stream},         {"stream_gen_polls", &Tuning::stream_gen_polls},
        {"stream_rows", &Tuning::stream_rows}, {"stream_ctas", &Tuning::stream_ctas},
        {"slabs", &Tuning::slabs},           {"no_slab", &Tuning::no_slab},
        {"stage", &Tuning::stage},           {"copy_threads", &Tuning::copy_threads}, {"populate", &Tuning::populate}, {"tile_smem_kb", &Tuning::tile_smem_kb}, {"tile_w", &Tuning::tile_w}, {"tile_wpc", &Tuning::tile_wpc},
        {"end_chunk", &Tuning::end_chunk},   {"pool", &Tuning::pool},
        {"pool_max", &Tuning::pool_max},     {"tile_direct", &Tuning::tile_direct},
        {"hugepage", &Tuning::hugepage},
    };
    std::lock_guard<std::mutex> lk(g_mu);
    if (strcmp(name, "reset") == 0) {
        g_tuning = Tuning{};
        return FRB_OK;
    }
    for (const Entry &e : table)
        if (strcmp(name, e.name) == 0) {
            g_tuning.*(e.field) = (int)value;
            return FRB_OK;
        }
    return fail(FRB_ERR_INVALID, "unknown tuning knob '%s'", name);
#else
    (void)value;
    return fail(FRB_ERR_INVALID, "this is a product build: tuning knob '%s' is not settable (build libfrb_tune.so with -DFRB_TUNING)", name);
#endif
}

void frb_shutdown(void)
{
    std::lock_guard<std::mutex> lk(g_mu);
    stop_workers_locked();
    g_copy.shutdown();
    g_stage.clear();
    for (int d = 0; d < kMaxDevices; ++d) {
        ctx_destroy(g_ctx[d]);
        g_ctx[d] = nullptr;
    }
    g_device_mask = 0;
    g_device_list.clear();
}

int frb_set_kernel_variant(uint32_t variant)
{
    if (variant > FRB_KERNEL_TILE || variant == 1u) return fail(FRB_ERR_INVALID, "unknown kernel variant %u", variant);
    tl_variant = variant;
    // also forgets the tile-kernel feedback (make_plan / harvest), so that a caller comparing kernels starts from a known state
    std::lock_guard<std::mutex> lk(g_mu);
    for (int d = 0; d < kMaxDevices; ++d)
        if (g_ctx[d]) g_ctx[d]->tile_bad_frames.store(0u);
    return FRB_OK;
}

int frb_render_shard(const FrbConfig *cfg, const FrbShard *shard, uint8_t *out, size_t out_len, uint32_t format, int device)
{
    if (!cfg) return fail(FRB_ERR_INVALID, "null config");
    if (format != FRB_RGB24 && format != FRB_RGBA32) return fail(FRB_ERR_INVALID, "unknown pixel format %u", format);
    const FrbShard s = shard ? *shard : whole_frame();
    return render_blocking(*cfg, s, device, format, out, out_len, nullptr, nullptr);
}

#ifdef FRB_STREAM_TRACE
// debug build only: globaltimer stamps of the last streamed render on `device` (see frb_stream.cuh)
extern "C" int frb_debug_stream_times(int device, unsigned long long *out16)
{
    DeviceCtx *c = get_ctx(device);
    if (!c) return tl_code;
    CU(cudaSetDevice(device));
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(out16, c->b().stats, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return FRB_OK;
}
extern "C" int frb_debug_warp_trace(unsigned int *trips, unsigned int *smid, unsigned long long *exit_ns)
{
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpyFromSymbol(trips, frb::g_warp_trips, sizeof(unsigned int) * 16384));
    CU(cudaMemcpyFromSymbol(smid, frb::g_warp_smid, sizeof(unsigned int) * 16384));
    CU(cudaMemcpyFromSymbol(exit_ns, frb::g_warp_exit, sizeof(unsigned long long) * 16384));
    return FRB_OK;
}
extern "C" int frb_debug_stream_trace(unsigned long long *out, unsigned int *n, int reset)
{
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpyFromSymbol(n, frb::g_trace_n, sizeof(unsigned int)));
    CU(cudaMemcpyFromSymbol(out, frb::g_trace, sizeof(unsigned long long) * 3 * 8192));
    if (reset) {
        const unsigned int z = 0;
        CU(cudaMemcpyToSymbol(frb::g_trace_n, &z, sizeof z));
    }
    return FRB_OK;
}
#endif

uint32_t frb_fern_default_threads(void)
{
    // rayon::current_num_threads() of the global pool: RAYON_NUM_THREADS if set to a positive number, else the logical CPUs
    if (const char *e = getenv("RAYON_NUM_THREADS")) {
        const long v = atol(e);
        if (v > 0) return (uint32_t)std::min<long>(v, 65536);
    }
    const unsigned hc = std::thread::hardware_concurrency();
    return hc ? hc : 1u;
}

int frb_render_fern(const FrbConfig *cfg, uint32_t threads, uint64_t seed, const FrbShard *shard, uint8_t *out, size_t out_len,
                    uint32_t format, int device)
{
    const double t0 = now_ms();
    if (!cfg) return fail(FRB_ERR_INVALID, "null config");
    if (format != FRB_RGB24 && format != FRB_RGBA32) return fail(FRB_ERR_INVALID, "unknown pixel format %u", format);
    const FrbShard sh = shard ? *shard : whole_frame();
    if (!shard_valid(sh)) return fail(FRB_ERR_INVALID, "bad shard: index %u of count %u", sh.index, sh.count);
    if (threads == 0) threads = frb_fern_default_threads();
    if (threads > 65536u) return fail(FRB_ERR_INVALID, "fern: %u virtual threads (max 65536)", threads);
    DeviceCtx *c = get_ctx(device < 0 ? 0 : device);
    if (!c) return tl_code;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->dev));

    frb::FernParams p;
    memset(&p, 0, sizeof p);
    p.width = cfg->width;
    p.height = cfg->height;
    p.threads = threads;
    p.per = cfg->iterations / threads;  // `config.iterations / threads` (src/lib.rs:286): the remainder is dropped
    {
        const uint64_t chains = std::min<uint64_t>(std::max<uint64_t>(((uint64_t)p.per + 511u) / 512u, 1u), 65536u);
        p.chains = (uint32_t)chains;
    }
    p.format = format;
    p.rows = shard_rows(cfg->height, sh);
    p.band_rows = shard_band_rows(cfg->height, sh);
    p.shard_index = sh.index;
    p.shard_count = sh.count;
    p.primary = (uint32_t)cfg->primary_color.r | ((uint32_t)cfg->primary_color.g << 8) | ((uint32_t)cfg->primary_color.b << 16);
    p.secondary = (uint32_t)cfg->secondary_color.r | ((uint32_t)cfg->secondary_color.g << 8) | ((uint32_t)cfg->secondary_color.b << 16);
    p.seed = seed;
    p.npix = (uint64_t)cfg->width * cfg->height;
    p.pos_re = cfg->pos.re;
    p.pos_im = cfg->pos.im;
    p.wf = (double)cfg->width;
    p.hf = (double)cfg->height;
    {
        // `65.0 * config.scale.re * config.height as f64 * 0.006` (src/lib.rs:425-426), left to right, one rounding each
        volatile double a = 65.0 * cfg->scale.re;
        volatile double b = a * (double)cfg->height;
        volatile double cx = b * 0.006;
        p.esx = cx;
        volatile double d = 37.0 * cfg->scale.im;
        volatile double e = d * (double)cfg->height;
        volatile double cy = e * 0.006;
        p.esy = cy;
    }
    p.amount = cfg->color_weight;

    const size_t out_bytes = (size_t)p.rows * cfg->width * format;
    if (out_len < out_bytes) return fail(FRB_ERR_INVALID, "output buffer too small: %zu < %zu bytes", out_len, out_bytes);
    if (out_bytes && !out) return fail(FRB_ERR_INVALID, "null output buffer");
    if (p.npix * threads > (1ull << 34))
        return fail(FRB_ERR_TOO_LARGE, "fern: %u images of %llu pixels exceed the hit-count budget (2^34 counters)", threads,
                    (unsigned long long)p.npix);
    FrbStats s = {};
    s.pixels = (uint64_t)p.rows * cfg->width;
    s.iterations = (uint64_t)p.per * threads;  // plotted points
    s.devices = 1;
    if (p.npix == 0 || out_bytes == 0) {
        tl_stats = s;
        return FRB_OK;
    }
    if (int rc = ensure(c->fern_counts, c->fern_counts_cap, (size_t)(p.npix * threads))) return rc;
    if (int rc = ensure(c->fern_out, c->fern_out_cap, out_bytes + 64)) return rc;
    if (!c->fern_lut) CU(cudaMalloc(&c->fern_lut, frb::kFernLut * sizeof(uint32_t)));
    for (auto &e : c->ev_fern)
        if (!e) CU(cudaEventCreate(&e));
    p.counts = c->fern_counts;
    p.lut = c->fern_lut;
    p.out = c->fern_out;

    cudaStream_t st = c->stream;
    CU(cudaEventRecord(c->ev_fern[0], st));
    CU(cudaMemsetAsync(p.counts, 0, (size_t)(p.npix * threads) * sizeof(uint32_t), st));
    frb::frb_fern_lut_kernel<<<1, 32, 0, st>>>(p);
    const uint64_t nchains = (uint64_t)threads * p.chains;
    if (p.per > 0)
        frb::frb_fern_chaos_kernel<<<(unsigned)((nchains + frb::kFernThreads - 1) / frb::kFernThreads), frb::kFernThreads, 0, st>>>(p);
    {
        const uint64_t n_out = (uint64_t)p.rows * cfg->width;
        const unsigned grid = (unsigned)std::min<uint64_t>((n_out + 255) / 256, (uint64_t)c->sm_count * 32);
        frb::frb_fern_compose_kernel<<<grid, 256, 0, st>>>(p);
    }
    CU(cudaEventRecord(c->ev_fern[1], st));
    CU(cudaMemcpyAsync(out, p.out, out_bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, c->ev_fern[0], c->ev_fern[1]));
    s.iterate_ms = ms;
    s.device_ms = ms;
    s.total_ms = (float)(now_ms() - t0);
    s.kernel_launches = p.per > 0 ? 3u : 2u;
    tl_stats = s;
    return FRB_OK;
}

int frb_render_counts(const FrbConfig *cfg, const FrbShard *shard, uint32_t *iters, double *zfinal, int device)
{
    if (!cfg) return fail(FRB_ERR_INVALID, "null config");
    const FrbShard s = shard ? *shard : whole_frame();
    if (!iters && (uint64_t)cfg->width * shard_rows(cfg->height, s) > 0) return fail(FRB_ERR_INVALID, "null iters buffer");
    return render_blocking(*cfg, s, device < 0 ? 0 : device, 0, nullptr, 0, iters, zfinal);
}

int frb_render_bands(const FrbConfig *cfg, const FrbShard *shard, uint8_t *frame, size_t frame_len, uint32_t format, int device)
{
    if (!cfg) return fail(FRB_ERR_INVALID, "null config");
    if (format != FRB_RGB24 && format != FRB_RGBA32) return fail(FRB_ERR_INVALID, "unknown pixel format %u", format);
    const FrbShard s = shard ? *shard : whole_frame();
    const size_t need = (size_t)cfg->width * cfg->height * format;
    if (frame_len < need) return fail(FRB_ERR_INVALID, "frame buffer too small: %zu < %zu bytes", frame_len, need);
    return render_blocking(*cfg, s, device < 0 ? 0 : device, format, frame, frame_len, nullptr, nullptr, true);
}

int frb_render(const FrbConfig *cfg, uint8_t *out, size_t out_len, uint32_t format, int device)
{
    if (!cfg) return fail(FRB_ERR_INVALID, "null config");
    if (format != FRB_RGB24 && format != FRB_RGBA32) return fail(FRB_ERR_INVALID, "unknown pixel format %u", format);
    const size_t need = (size_t)cfg->width * cfg->height * format;
    if (out_len < need) return fail(FRB_ERR_INVALID, "output buffer too small: %zu < %zu bytes", out_len, need);
    if (device >= 0) return render_blocking(*cfg, whole_frame(), device, format, out, out_len, nullptr, nullptr);

    // device = -1: interleaved 16-row bands over all initialised devices, every device writing (pinned destination) or copying
    // its bands straight into the caller's image.  No collective: shards are independent.
    const std::vector<int> devs = active_devices();
    if (devs.empty()) return fail(FRB_ERR_NO_DEVICE, "no usable CUDA device; this library has no CPU path");
    if (devs.size() == 1 || cfg->height <= kDefaultBandRows)
        return render_blocking(*cfg, whole_frame(), devs[0], format, out, out_len, nullptr, nullptr);

    const double t0 = now_ms();
    const uint32_t G = (uint32_t)devs.size();
    struct Part {
        int rc = FRB_OK;
        std::string err;
        FrbStats stats = {};
    };
    std::vector<Part> parts(G);
    const uint32_t variant = tl_variant;  // the caller's kernel override travels to the workers
    // shard g on device devs[g]: plan, enqueue, wait, harvest — on that device's own host thread
    auto run_part = [&](uint32_t g) {
        tl_variant = variant;
        const FrbShard sh = {g, G, kDefaultBandRows, 0u};
        parts[g].rc = render_blocking(*cfg, sh, devs[g], format, out, out_len, nullptr, nullptr, true);
        if (parts[g].rc != FRB_OK) parts[g].err = tl_error;
        else parts[g].stats = tl_stats;
    };
    // one worker per SHARD slot: a device listed twice gets two (they serialise on its mutex)
    run_parts_on_workers(G, run_part);
    FrbStats total = {};
    for (uint32_t g = 0; g < G; ++g) {
        if (parts[g].rc != FRB_OK) return fail(parts[g].rc, "device %d: %s", devs[g], parts[g].err.c_str());
        const FrbStats &s = parts[g].stats;
        total.pixels += s.pixels;
        total.iterations += s.iterations;
        total.lane_iterations += s.lane_iterations;
        total.iterate_ms = std::max(total.iterate_ms, s.iterate_ms);
        total.shade_ms = std::max(total.shade_ms, s.shade_ms);
        total.device_ms = std::max(total.device_ms, s.device_ms);
        total.kernel_launches += s.kernel_launches;
        total.kernel_variant = s.kernel_variant;
    }
    total.devices = G;
    total.total_ms = (float)(now_ms() - t0);
    tl_stats = total;
    return FRB_OK;
}

int frb_render_frames(const FrbConfig *cfgs, uint32_t n, uint8_t *out, size_t out_len, uint32_t format, int device)
{
    return frb_render_frames_strided(cfgs, n, out, out_len, format, device, 0u, 1u);
}

int frb_render_frames_strided(const FrbConfig *cfgs, uint32_t n, uint8_t *out, size_t out_len, uint32_t format, int device,
                              uint32_t first, uint32_t stride)
{
    if (!cfgs && n) return fail(FRB_ERR_INVALID, "null configs");
    if (stride == 0u) return fail(FRB_ERR_INVALID, "zero frame stride");
    const uint32_t mine = first < n ? (n - first + stride - 1u) / stride : 0u;  // frames first, first + stride, ... below n
    if (format != FRB_RGB24 && format != FRB_RGBA32) return fail(FRB_ERR_INVALID, "unknown pixel format %u", format);
    size_t need = 0;
    for (uint32_t k = 0; k < n; ++k) need += (size_t)cfgs[k].width * cfgs[k].height * format;
    if (out_len < need) return fail(FRB_ERR_INVALID, "output buffer too small: %zu < %zu bytes", out_len, need);
    std::vector<int> devs;
    if (device >= 0)
        devs.push_back(device);
    else
        devs = active_devices();
    // one worker per DEVICE: a device listed twice (frb_init_list) still renders its frames on its two buffer sets only
    std::sort(devs.begin(), devs.end());
    devs.erase(std::unique(devs.begin(), devs.end()), devs.end());
    if (devs.empty()) return fail(FRB_ERR_NO_DEVICE, "no usable CUDA device; this library has no CPU path");
    const double t0 = now_ms();
    const uint32_t G = (uint32_t)devs.size();
    std::vector<size_t> offset(n);
    {
        size_t off = 0;
        for (uint32_t k = 0; k < n; ++k) {
            offset[k] = off;
            off += (size_t)cfgs[k].width * cfgs[k].height * format;
        }
    }
    // Frames are independent and differ 80x in cost over a zoom, so they are handed out IN ORDER from one atomic
    // counter to whichever device is ready for another (imbalance 1.011 against 1.018 for static round robin, 2.63 for blocks).
    // Every device has its own host thread that keeps two frames in flight on alternating buffer sets and streams: frame j+1's
    // persistent CTAs start as frame j's drain, and frame j's shade + copy-out run in frame j+1's shadow.  The device's mutex
    // is held per frame (enqueue, harvest), not per call: a get_image from another thread — the GUI's screenshot thread,
    // src/gui.rs:325 — slips in between two frames of an animation instead of waiting for all of it.
    std::atomic<uint32_t> next{0};
    std::atomic<bool> failed{false};
    struct Part {
        int rc = FRB_OK;
        std::string err;
        FrbStats stats = {};
    };
    std::vector<Part> parts(G);
    const uint32_t variant = tl_variant;
    auto run_device = [&](uint32_t g) -> int {
        DeviceCtx *cp = get_ctx(devs[g]);
        if (!cp) return tl_code;
        DeviceCtx &c = *cp;
        CU(cudaSetDevice(c.dev));
        FrbStats &total = parts[g].stats;
        struct InFlight {
            Slot *slot;
            cudaEvent_t done;
        };
        InFlight pending[2] = {{nullptr, nullptr}, {nullptr, nullptr}};
        int npending = 0;
        uint32_t nth = 0;  // frames this device has issued
        auto retire_oldest = [&]() -> int {
            const InFlight f = pending[0];
            pending[0] = pending[1];
            --npending;
            CU(cudaEventSynchronize(f.done));  // not under the mutex: other callers may use the device meanwhile
            std::lock_guard<std::mutex> lk(c.mu);
            FrbStats st = {};
            if (int rc = harvest(c, *f.slot, &st, 0.0)) return rc;
            total.pixels += st.pixels;
            total.iterations += st.iterations;
            total.lane_iterations += st.lane_iterations;
            total.iterate_ms += st.iterate_ms;
            total.shade_ms += st.shade_ms;
            total.device_ms += st.device_ms;
            total.kernel_launches += st.kernel_launches;
            total.kernel_variant = st.kernel_variant;
            return FRB_OK;
        };
        for (;;) {
            while (npending < 2 && !failed.load(std::memory_order_relaxed)) {
                const uint32_t j = next.fetch_add(1, std::memory_order_relaxed);
                if (j >= mine) break;
                const uint32_t k = first + j * stride;
                // a blocking render from another thread goes first (bounded: it holds the device for one frame of its own)
                for (int spin = 0; c.waiting.load(std::memory_order_relaxed) > 0 && spin < 4000; ++spin)
                    std::this_thread::sleep_for(std::chrono::microseconds(50));
                std::lock_guard<std::mutex> lk(c.mu);
                const int alt = (int)(nth++ & 1u);
                c.cur_set = alt;
                cudaStream_t st = c.aux[alt];
                Plan plan;
                int rc = make_plan(c, cfgs[k], whole_frame(), format, &plan);
                if (!rc) rc = ensure(c.b().out, c.b().out_cap, plan.out_bytes + 64);
                Slot *slot = nullptr;
                if (!rc) rc = enqueue(c, plan, c.b().out, st, &slot);
                if (!rc && plan.out_bytes && cudaMemcpyAsync(out + offset[k], c.b().out, plan.out_bytes, cudaMemcpyDeviceToHost, st) != cudaSuccess)
                    rc = fail(FRB_ERR_CUDA, "cudaMemcpyAsync (frame %u) failed: %s", k, cudaGetErrorString(cudaGetLastError()));
                if (!rc) rc = finish_enqueue(c, *slot, st);
                c.cur_set = 0;
                if (rc) {
                    drain_streams(c);  // up to two frames' copies into the caller's `out` are still in flight
                    return rc;
                }
                pending[npending].slot = slot;
                pending[npending].done = slot->ev_end;
                ++npending;
            }
            if (npending == 0) break;
            if (int rc = retire_oldest()) {
                drain_streams(c);
                return rc;
            }
        }
        return FRB_OK;
    };
    auto run_part = [&](uint32_t g) {
        tl_variant = variant;
        parts[g].rc = run_device(g);
        if (parts[g].rc != FRB_OK) {
            parts[g].err = tl_error;
            failed.store(true);
        }
    };
    run_parts_on_workers(G, run_part);
    FrbStats total = {};
    for (uint32_t g = 0; g < G; ++g) {
        if (parts[g].rc != FRB_OK) return fail(parts[g].rc, "device %d: %s", devs[g], parts[g].err.c_str());
        const FrbStats &s = parts[g].stats;
        total.pixels += s.pixels;
        total.iterations += s.iterations;
        total.lane_iterations += s.lane_iterations;
        total.iterate_ms += s.iterate_ms;
        total.shade_ms += s.shade_ms;
        total.device_ms = std::max(total.device_ms, s.device_ms);  // the busiest device's total: the critical path
        total.kernel_launches += s.kernel_launches;
        total.kernel_variant = s.kernel_variant;
    }
    total.devices = G;
    total.total_ms = (float)(now_ms() - t0);
    tl_stats = total;
    return FRB_OK;
}

int frb_render_device_async(const FrbConfig *cfg, const FrbShard *shard, void *dev_out, size_t out_len, uint32_t format, int device,
                            void *stream)
{
    if (!cfg) return fail(FRB_ERR_INVALID, "null config");
    if (format != FRB_RGB24 && format != FRB_RGBA32) return fail(FRB_ERR_INVALID, "unknown pixel format %u", format);
    const FrbShard s = shard ? *shard : whole_frame();
    DeviceCtx *c = get_ctx(device);
    if (!c) return tl_code;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(device));
    cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
    // Renders issued on two different caller streams use the two sets of device buffers (tables, records, queue, statistics),
    // so they can overlap like the frames of frb_render_frames: the next frame's persistent CTAs start on every SM the moment
    // the previous frame's CTA leaves it, and its shade kernel runs beside them.  A stream keeps the set it used last; a stream
    // that owns neither takes the set that has been idle longer — so two caller streams end up on different sets whatever ran
    // before them (a blocking render in between leaves set 0 with the library's stream: "the other one" then put BOTH caller
    // streams on set 1, one frame at a time).  The library's own stream: set 0, strictly in order, as before.
    int set = 0;
    if (stream != nullptr) {
        if (c->sets[1].have_guard && c->sets[1].last_stream == st) set = 1;
        else if (c->sets[0].have_guard && c->sets[0].last_stream == st) set = 0;
        else if (c->sets[0].have_guard && (!c->sets[1].have_guard || c->sets[1].last_use < c->sets[0].last_use)) set = 1;
    }
    struct SetGuard {
        DeviceCtx *c;
        ~SetGuard() { c->cur_set = 0; }
    } set_guard{c};
    c->cur_set = set;
    Plan plan;
    if (int rc = make_plan(*c, *cfg, s, format, &plan)) return rc;
    if (out_len < plan.out_bytes) return fail(FRB_ERR_INVALID, "output buffer too small: %zu < %zu bytes", out_len, plan.out_bytes);
    if (plan.out_bytes && !dev_out) return fail(FRB_ERR_INVALID, "null device output buffer");
    if (((uintptr_t)dev_out & 15u) != 0) return fail(FRB_ERR_INVALID, "device output buffer must be 16-byte aligned");
    const double t0 = now_ms();
    Slot *slot = nullptr;
    if (int rc = enqueue(*c, plan, (uint8_t *)dev_out, st, &slot)) return rc;
    if (int rc = finish_enqueue(*c, *slot, st)) return rc;
    AsyncRecord &rec = tl_async_ring[tl_async_count % kSlots];
    rec.dev = device;
    rec.slot = slot;
    rec.t0 = t0;
    ++tl_async_count;
    return FRB_OK;
}

int frb_async_stats(uint32_t back, FrbStats *stats)
{
    if (tl_async_count == 0) return fail(FRB_ERR_INVALID, "no asynchronous render issued on this thread");
    if (back >= (uint32_t)kSlots / 2 || back >= tl_async_count)
        return fail(FRB_ERR_INVALID, "statistics of the render %u calls ago are no longer kept (ring of %d)", back, kSlots / 2);
    const AsyncRecord &rec = tl_async_ring[(tl_async_count - 1 - back) % kSlots];
    DeviceCtx *c = get_ctx(rec.dev);
    if (!c) return tl_code;
    CU(cudaSetDevice(rec.dev));
    FrbStats s;
    if (int rc = harvest(*c, *rec.slot, &s, now_ms() - rec.t0)) return rc;
    tl_stats = s;
    if (stats) *stats = s;
    return FRB_OK;
}

int frb_sync_stats(FrbStats *stats) { return frb_async_stats(0, stats); }

int frb_last_stats(FrbStats *stats)
{
    if (!stats) return fail(FRB_ERR_INVALID, "null stats");
    *stats = tl_stats;
    return FRB_OK;
}

}  // extern "C"
