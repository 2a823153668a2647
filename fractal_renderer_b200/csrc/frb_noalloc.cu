// frb_noalloc.cu — frb_render_static_async: a whole frame into the CALLER's device buffer with no allocation of any kind.
//
// Who this is for.  Every other render entry point of libfrb grows the library's own device buffers on demand (cudaMalloc in
// make_plan / ctx_init of frb_api.cu) and keeps pinned host slots for its statistics.  Three kinds of caller cannot live with
// that:
//   * a host whose GPU memory has been handed out in full before we are called — a frame server that lives inside a process
//     whose own pool allocator took (nearly) all of HBM at start-up, so that a cudaMalloc issued from a plug-in fails or, worse,
//     succeeds by pushing the host's next allocation over the edge;
//   * a render loop with a latency budget: cudaMalloc and cudaFree synchronise the whole device, so the first frame of every
//     new size stalls whatever else the process has in flight — an interactive viewer (the reference's GUI thread,
//     src/gui.rs:56-82) resizes on every drag of the window edge;
//   * code that runs where the CUDA runtime's allocator must not be entered at all — inside another library's callback, under a
//     lock that its allocator also takes, or in a sandbox whose policy allows launches but no new device mappings.
// For them the scratch memory of the frame — coordinate tables, the raw (count, final z) records between the iterate and the
// shade kernel, the chunk counter and the statistics words — lives in ONE file-scope `static __device__` array of fixed size,
// which the CUDA loader places once per device when this library's code is loaded there.  The price is a hard ceiling on the
// frame size (kStaticMaxPixels, kStaticMaxEdgeSum below) and ~20 MiB of device memory that every GPU running code from this file
// holds until the process exits, whether or not the entry point is ever called.
//
// The kernels are the library's own (K0 tables -> K1 persistent lane-refill iterate, 7-instruction form: exact for ANY input
// -> K2 shade; frb_kernels.cuh), compiled a second time into this translation unit under another namespace, because a
// `static __device__` array is only nameable — by device code directly, by host code through cudaGetSymbolAddress — inside
// the .cu file that defines it.  NB in HOST code the identifier g_static_scratch is a host-side shadow object, not device
// memory: handing `g_static_scratch` itself to a kernel would fault (or, on a system where the GPU can reach host memory,
// quietly scribble over the wrong bytes).  static_setup() asks the runtime for the real address, once per device, with that
// device current, and only that pointer is ever passed on.
//
// Nothing here replaces a reference function (the reference allocates a fresh Vec per call, src/lib.rs:266); the path computed
// is the same get_image Mandelbrot|Julia arm (src/lib.rs:253-270, calc/src/lib.rs:199-257) as frb_render_device_async.
#include "../../include/frb.h"

#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

// This file's private copies of the kernels: same source, namespace frb_static instead of frb, so that the host-side launch
// stubs do not collide with the ones frb_api.cu defines.
#define frb frb_static
#include "frb_kernels.cuh"
#undef frb

int frb_internal_fail(int code, const char *fmt, ...);  // frb_api.cu: sets frb_last_error() of the calling thread

namespace {

using frb_static::RenderParams;

// ---- the ceiling ------------------------------------------------------------------------------------------------------------
constexpr uint32_t kStaticMaxPixels = 1u << 20;    // width * height  (1024 x 1024, 1280 x 720, 1366 x 767, ...)
constexpr uint32_t kStaticMaxEdgeSum = 16384;      // width + height  (the two coordinate tables)

// ---- layout of the array (every part starts on a 256-byte boundary) ---------------------------------------------------------
constexpr size_t kStatsOff = 0;                                             // 128 x u64: statistics + control words (K0 zeroes what it uses)
constexpr size_t kQueueOff = 1024;                                          // 64 x u32: chunk counter(s) (K0 zeroes them)
constexpr size_t kTabOff = 1280;                                            // (width + height) x f64, written by K0 before it is read
constexpr size_t kItersOff = kTabOff + (size_t)kStaticMaxEdgeSum * 8;       // pixels x u32, every word written by K1 before K2 reads it
constexpr size_t kZOff = kItersOff + (size_t)kStaticMaxPixels * 4;          // pixels x 2 f64, likewise
constexpr size_t kStaticScratchBytes = kZOff + (size_t)kStaticMaxPixels * 16;  // 21 103 872 bytes = 20.1 MiB
static_assert(kTabOff % 256 == 0 && kItersOff % 256 == 0 && kZOff % 256 == 0, "parts of the static scratch are 256-byte aligned");

// Device memory, one instance per GPU that loads this code, zero-filled by the loader, alive until the process ends.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 32;
struct StaticDevice {
    unsigned char *scratch = nullptr;  // DEVICE address of g_static_scratch on this GPU (cudaGetSymbolAddress), never the host symbol
    int sm_count = 0;
    int iterate_ctas_per_sm = 0;
    bool ready = false;
};
StaticDevice g_static_dev[kMaxDevices];
std::mutex g_static_mu;

#define CU_STATIC(call)                                                                                                    \
    do {                                                                                                                   \
        cudaError_t e_ = (call);                                                                                           \
        if (e_ != cudaSuccess)                                                                                             \
            return frb_internal_fail(FRB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// Once per device, with that device current (the caller has just called cudaSetDevice): where does this GPU keep the array,
// how many CTAs of the iterate kernel fit.  Also makes the runtime load the three kernels' code now rather than at their first
// launch.  Allocates nothing itself.
int static_setup(int device, const StaticDevice **out)
{
    std::lock_guard<std::mutex> lk(g_static_mu);
    StaticDevice &d = g_static_dev[device];
    if (!d.ready) {
        int major = 0;
        CU_STATIC(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
        if (major < 10) return frb_internal_fail(FRB_ERR_NO_DEVICE, "device %d is not sm_100: this library carries sm_100a code only", device);
        void *ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (e != cudaSuccess || ptr == nullptr)
            return frb_internal_fail(FRB_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", device, cudaGetErrorString(e));
        if (((uintptr_t)ptr & 255u) != 0) return frb_internal_fail(FRB_ERR_CUDA, "static scratch of device %d is not 256-byte aligned", device);
        CU_STATIC(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device));
        CU_STATIC(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.iterate_ctas_per_sm, frb_static::frb_iterate_stream_kernel<7>,
                                                                frb_static::kIterateThreads, 0));
        if (d.sm_count < 1 || d.iterate_ctas_per_sm < 1) return frb_internal_fail(FRB_ERR_CUDA, "iterate kernel does not fit on device %d", device);
        cudaFuncAttributes attr;
        CU_STATIC(cudaFuncGetAttributes(&attr, frb_static::frb_coord_tables_kernel));
        CU_STATIC(cudaFuncGetAttributes(&attr, frb_static::frb_iterate_stream_kernel<7>));
        CU_STATIC(cudaFuncGetAttributes(&attr, frb_static::frb_shade_kernel));
        d.scratch = static_cast<unsigned char *>(ptr);
        d.ready = true;
    }
    *out = &d;
    return FRB_OK;
}

uint64_t bits_of(double v)
{
    uint64_t b;
    memcpy(&b, &v, sizeof b);
    return b;
}

}  // namespace

extern "C" int frb_render_static_async(const FrbConfig *cfg, void *dev_out, size_t out_len, uint32_t format, int device, void *stream)
{
    if (!cfg) return frb_internal_fail(FRB_ERR_INVALID, "null config");
    if (format != FRB_RGB24 && format != FRB_RGBA32) return frb_internal_fail(FRB_ERR_INVALID, "unknown pixel format %u", format);
    if (cfg->algo > FRB_ALGO_JULIA) return frb_internal_fail(FRB_ERR_INVALID, "unknown algo %u", cfg->algo);
    const uint64_t npix = (uint64_t)cfg->width * cfg->height;
    if (npix > kStaticMaxPixels || (uint64_t)cfg->width + cfg->height > kStaticMaxEdgeSum)
        return frb_internal_fail(FRB_ERR_TOO_LARGE,
                                 "frb_render_static_async: %u x %u exceeds the static scratch (width * height <= %u and width + height <= %u); "
                                 "larger frames need an allocating entry point (frb_render_device_async)",
                                 cfg->width, cfg->height, kStaticMaxPixels, kStaticMaxEdgeSum);
    const size_t out_bytes = (size_t)npix * format;
    if (out_len < out_bytes) return frb_internal_fail(FRB_ERR_INVALID, "output buffer too small: %zu < %zu bytes", out_len, out_bytes);
    if (out_bytes && !dev_out) return frb_internal_fail(FRB_ERR_INVALID, "null device output buffer");
    if (((uintptr_t)dev_out & 15u) != 0) return frb_internal_fail(FRB_ERR_INVALID, "device output buffer must be 16-byte aligned");
    const int n = frb_device_count();
    if (n == 0) return frb_internal_fail(FRB_ERR_NO_DEVICE, "no usable CUDA device (cudaGetDeviceCount found none); this library has no CPU path");
    if (device < 0 || device >= n || device >= kMaxDevices) return frb_internal_fail(FRB_ERR_INVALID, "device %d out of range [0, %d)", device, n);
    CU_STATIC(cudaSetDevice(device));
    const StaticDevice *d = nullptr;
    if (int rc = static_setup(device, &d)) return rc;
    if (npix == 0) return FRB_OK;  // an empty image: nothing to write

    // The frame's parameters, as make_plan (frb_api.cu) fills them for a whole frame and the STREAM7 kernel.
    RenderParams p;
    memset(&p, 0, sizeof p);
    p.width = cfg->width;
    p.height = cfg->height;
    p.rows = cfg->height;
    p.band_rows = std::max(cfg->height, 1u);
    p.band_shift = 32u;
    p.shard_index = 0;
    p.shard_count = 1;
    p.iterations = cfg->iterations;
    p.algo = cfg->algo;
    {
        volatile double l = cfg->limit;  // `let squared = limit * limit` (calc/src/lib.rs:246): one IEEE multiply
        volatile double s = l * l;
        p.limit_sq = s;
    }
    p.filter_hi = (uint32_t)(bits_of(p.limit_sq * 0.5) >> 32);
    p.chunk = 64u;
    p.chunks_per_row = (cfg->width + p.chunk - 1) / p.chunk;
    p.total_chunks = p.rows * p.chunks_per_row;  // <= 2^20 rows x chunks: no overflow under the ceiling
    p.julia_re = cfg->julia_set.re;
    p.julia_im = cfg->julia_set.im;
    p.pos_re = cfg->pos.re;
    p.pos_im = cfg->pos.im;
    p.scale_re = cfg->scale.re;
    p.scale_im = cfg->scale.im;
    p.stable_limit = cfg->stable_limit;
    p.exposure = cfg->exposure;
    p.smooth = cfg->smooth ? 1u : 0u;
    p.inside = cfg->inside ? 1u : 0u;
    p.primary = (uint32_t)cfg->primary_color.r | ((uint32_t)cfg->primary_color.g << 8) | ((uint32_t)cfg->primary_color.b << 16);
    p.secondary = (uint32_t)cfg->secondary_color.r | ((uint32_t)cfg->secondary_color.g << 8) | ((uint32_t)cfg->secondary_color.b << 16);
    p.format = format;
    p.iterations_d = (double)cfg->iterations;
    p.inv_iterations = cfg->iterations ? 1.0 / p.iterations_d : 0.0;
    p.prim_r = (double)cfg->primary_color.r;
    p.prim_g = (double)cfg->primary_color.g;
    p.prim_b = (double)cfg->primary_color.b;
    p.fast_shade = (cfg->iterations >= 1u && std::isfinite(cfg->exposure) && std::fabs(cfg->exposure) < 1e290) ? 1u : 0u;
    // every buffer but the image is a slice of THIS device's copy of the static array
    unsigned char *const base = d->scratch;
    p.stats = reinterpret_cast<unsigned long long *>(base + kStatsOff);
    p.queue = reinterpret_cast<unsigned int *>(base + kQueueOff);
    p.tab_re = reinterpret_cast<double *>(base + kTabOff);
    p.tab_im = p.tab_re + cfg->width;
    p.iters = reinterpret_cast<uint32_t *>(base + kItersOff);
    p.zfinal = reinterpret_cast<double2 *>(base + kZOff);
    p.out = static_cast<uint8_t *>(dev_out);

    cudaStream_t st = static_cast<cudaStream_t>(stream);  // NULL = the legacy default stream: creating one of our own would allocate
    // K0 writes the tables and ZEROES the words that must start at zero (chunk counter, statistics): the array still holds the
    // previous call's contents at this point.
    const uint32_t edge = std::max(p.width, p.height);
    frb_static::frb_coord_tables_kernel<<<(edge + 255) / 256, 256, 0, st>>>(p);
    if (cfg->algo == FRB_ALGO_BARNSLEY_FERN) {
        // get_recursive_pixel returns BLACK without iterating (calc/src/lib.rs:211): zero records shade to black
        CU_STATIC(cudaMemsetAsync(p.iters, 0, npix * sizeof(uint32_t), st));
        CU_STATIC(cudaMemsetAsync(p.zfinal, 0, npix * sizeof(double2), st));
    } else {
        const uint32_t warps_per_cta = frb_static::kIterateThreads / 32;
        const uint32_t needed = (p.total_chunks + warps_per_cta - 1) / warps_per_cta;
        const uint32_t ctas = std::max(1u, std::min((uint32_t)(d->sm_count * d->iterate_ctas_per_sm), needed));
        frb_static::frb_iterate_stream_kernel<7><<<ctas, frb_static::kIterateThreads, 0, st>>>(p);
    }
    const uint64_t cta_px = (uint64_t)frb_static::kShadeThreads * frb_static::kShadePxPerThread;
    const uint32_t shade_ctas = (uint32_t)std::min<uint64_t>((npix + cta_px - 1) / cta_px, (uint64_t)d->sm_count * 32);
    frb_static::frb_shade_kernel<<<shade_ctas, frb_static::kShadeThreads, 0, st>>>(p);
    CU_STATIC(cudaGetLastError());
    return FRB_OK;
}
