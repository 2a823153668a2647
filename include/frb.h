/*
 * frb.h — C ABI of the B200-native escape-time renderer ("frb" = fractal renderer, B200).
 *
 * This is the drop-in boundary for ONE path of Icelk/fractal-renderer: the Mandelbrot | Julia arm of
 *     pub fn get_image(config: &Config) -> Vec<RGB>            (reference src/lib.rs:253-270)
 * i.e. calc::get_recursive_pixel (calc/src/lib.rs:199-235) evaluated for every pixel, row-major.
 * Everything behind these symbols runs as hand-written sm_100a CUDA; there is NO CPU fallback: every
 * compute entry point returns FRB_ERR_NO_DEVICE when no CUDA device is usable.
 *
 * Conventions
 *   - plain C types only; the caller owns every host buffer; the library owns device memory, pinned staging
 *     and streams (created lazily or by frb_init, released by frb_shutdown);
 *   - every function returns 0 (FRB_OK) or a negative FRB_ERR_* code, never unwinds; the message for the last
 *     failure on the calling thread is frb_last_error();
 *   - calls are blocking unless the name ends in _async; the library is thread-safe (concurrent renders from
 *     several host threads are serialised per device — reference callers src/gui.rs:60 and :325 may overlap);
 *   - images are row-major, y outer, no row padding; RGB24 = 3 bytes r,g,b per pixel (the memory image of the
 *     reference's Vec<RGB>, src/lib.rs:13-15), RGBA32 = r,g,b,255 (what src/gui.rs:66-75 builds for egui).
 */
#ifndef FRB_H_
#define FRB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FRB_ABI_VERSION 1u

/* ---- value types: field-for-field mirrors of the reference's calc crate --------------------------------- */

/* calc::Imaginary (calc/src/lib.rs:77-81) */
typedef struct FrbImaginary {
    double re;
    double im;
} FrbImaginary;

/* calc::RGB (calc/src/lib.rs:119-125): the three STORED fields in declaration order r,g,b.
 * NB the reference constructor is RGB::new(r, b, g) (calc/src/lib.rs:129-131): callers that mirror
 * `RGB::new(40, 40, 255)` must store r=40, g=255, b=40. frb_rgb_new() below does that. */
typedef struct FrbRGB {
    uint8_t r;
    uint8_t g;
    uint8_t b;
} FrbRGB;

/* calc::Algo (calc/src/lib.rs:150-154), declaration order. */
enum {
    FRB_ALGO_MANDELBROT = 0,
    FRB_ALGO_BARNSLEY_FERN = 1, /* frb_render*: all-black, as get_recursive_pixel does (:211); the chaos game is frb_render_fern */
    FRB_ALGO_JULIA = 2
};

/* calc::Config (calc/src/lib.rs:21-37), same field order; bools as bytes. 104 bytes, 8-aligned. */
typedef struct FrbConfig {
    uint32_t algo;        /* FRB_ALGO_* */
    uint32_t width;
    uint32_t height;
    uint32_t iterations;
    double limit;         /* escape radius (un-squared) */
    double stable_limit;  /* compared with the SQUARED distance of the final z (calc/src/lib.rs:216, sic) */
    FrbImaginary pos;
    FrbImaginary scale;
    double exposure;
    uint8_t inside;
    uint8_t smooth;
    FrbRGB primary_color;
    FrbRGB secondary_color;
    double color_weight;  /* fern only (frb_render_fern): subtract_pixel's `amount` */
    FrbImaginary julia_set;
} FrbConfig;

/* Layout checks: a binding that mirrors these structs (#[repr(C)] in rust/ffi.rs, ctypes in the Python package) must agree with
 * every offset below; a mismatch is a compile error on the C side too. */
#if defined(__cplusplus) && __cplusplus >= 201103L
#define FRB_STATIC_ASSERT(cond, msg) static_assert(cond, msg)
#elif defined(__STDC_VERSION__) && __STDC_VERSION__ >= 201112L
#define FRB_STATIC_ASSERT(cond, msg) _Static_assert(cond, msg)
#else /* C89 / C99: a negative array size is the classic compile-time check */
#define FRB_SA_CAT2(a, b) a##b
#define FRB_SA_CAT(a, b) FRB_SA_CAT2(a, b)
#define FRB_STATIC_ASSERT(cond, msg) typedef char FRB_SA_CAT(frb_static_assert_line_, __LINE__)[(cond) ? 1 : -1]
#endif
FRB_STATIC_ASSERT(sizeof(FrbImaginary) == 16, "FrbImaginary is two doubles");
FRB_STATIC_ASSERT(sizeof(FrbRGB) == 3, "FrbRGB is three packed bytes (the memory image of calc::RGB)");
FRB_STATIC_ASSERT(sizeof(FrbConfig) == 104, "FrbConfig is 104 bytes");
FRB_STATIC_ASSERT(offsetof(FrbConfig, algo) == 0 && offsetof(FrbConfig, width) == 4 && offsetof(FrbConfig, height) == 8 &&
                      offsetof(FrbConfig, iterations) == 12,
                  "FrbConfig: algo, width, height, iterations");
FRB_STATIC_ASSERT(offsetof(FrbConfig, limit) == 16 && offsetof(FrbConfig, stable_limit) == 24, "FrbConfig: limit, stable_limit");
FRB_STATIC_ASSERT(offsetof(FrbConfig, pos) == 32 && offsetof(FrbConfig, scale) == 48 && offsetof(FrbConfig, exposure) == 64,
                  "FrbConfig: pos, scale, exposure");
FRB_STATIC_ASSERT(offsetof(FrbConfig, inside) == 72 && offsetof(FrbConfig, smooth) == 73 && offsetof(FrbConfig, primary_color) == 74 &&
                      offsetof(FrbConfig, secondary_color) == 77,
                  "FrbConfig: inside, smooth, primary_color, secondary_color");
FRB_STATIC_ASSERT(offsetof(FrbConfig, color_weight) == 80 && offsetof(FrbConfig, julia_set) == 88, "FrbConfig: color_weight, julia_set");

/* Fills *cfg with calc::Config::new(algo) (calc/src/lib.rs:39-69). Host-only, needs no device. */
void frb_config_default(FrbConfig *cfg, uint32_t algo);
/* RGB::new(r, b, g) with the reference's argument order (calc/src/lib.rs:129-131). Host-only. */
FrbRGB frb_rgb_new(uint8_t r, uint8_t b, uint8_t g);

/* ---- pixel formats / errors -------------------------------------------------------------------------------- */

enum { FRB_RGB24 = 3, FRB_RGBA32 = 4 }; /* value = bytes per pixel */

enum {
    FRB_OK = 0,
    FRB_ERR_NO_DEVICE = -1,   /* no usable CUDA device / driver: there is no CPU path */
    FRB_ERR_INVALID = -2,     /* bad argument (null pointer, short buffer, unknown format, bad shard) */
    FRB_ERR_CUDA = -3,        /* a CUDA runtime call failed; see frb_last_error() */
    FRB_ERR_TOO_LARGE = -4    /* a single launch would exceed 2^32-1 pixels per shard */
};

/* ---- shards (multi-GPU partition of one frame) ---------------------------------------------- */

/* Interleaved row bands: band k covers image rows [k*band_rows, (k+1)*band_rows); shard `index` of `count`
 * owns the bands with k % count == index. count=1,index=0,band_rows=0 means the whole frame.
 * A shard's output is COMPACT: its rows in increasing y, no gaps (frb_shard_rows() of them). */
typedef struct FrbShard {
    uint32_t index;
    uint32_t count;
    uint32_t band_rows; /* 0 = whole frame as one band */
    uint32_t reserved;
} FrbShard;

/* Number of image rows shard owns for an image of `height` rows. Host-only. */
uint32_t frb_shard_rows(uint32_t height, const FrbShard *shard);
/* Image row of the shard's local row `local_row` (image of `height` rows). Host-only. */
uint32_t frb_shard_row_to_y(uint32_t height, const FrbShard *shard, uint32_t local_row);

/* ---- statistics of the last render issued by the calling thread --------------------------------------- */

typedef struct FrbStats {
    uint64_t pixels;          /* pixels rendered */
    uint64_t iterations;      /* loop-body executions of calc::recursive summed over pixels (i+1 | iterations) */
    uint64_t lane_iterations; /* 32 x warp-level loop trips issued by the iterate kernel (lane utilisation = iterations / lane_iterations) */
    float iterate_ms;         /* device time of the iterate kernel(s), CUDA events */
    float shade_ms;           /* device time of the shade kernel(s) */
    float device_ms;          /* first kernel start -> last kernel end on the device */
    float total_ms;           /* host wall time of the call (includes copies for host-buffer entry points) */
    uint32_t kernel_launches; /* kernels launched by this render */
    uint32_t devices;         /* devices that took part */
    uint32_t kernel_variant;  /* FRB_KERNEL_* actually used for the iteration */
    uint32_t reserved;
} FrbStats;

enum {
    FRB_KERNEL_AUTO = 0,      /* TILE for short iteration caps and image formats; else DEFER6 when its preconditions hold (every
                                 sane input), else DENSE6, else STREAM7 */
    /* 1 was STREAM6 (positions of the last block recorded in registers), retired in favour of DEFER6 */
    FRB_KERNEL_STREAM7 = 2,   /* persistent lane-refill kernel, 7 FP64 instructions / iteration, filter every
                                 iteration: exact for ANY input */
    FRB_KERNEL_SIMPLE = 3,    /* one thread per pixel, reference operation order as written (11 FP64 instructions), no
                                 queue: the on-GPU cross-check for the others */
    FRB_KERNEL_DENSE6 = 4,    /* persistent lane-refill kernel, 6 FP64 instructions / iteration, filter every iteration
                                 (small or huge limits, iteration caps < 32) */
    FRB_KERNEL_DEFER6 = 5,    /* the default: 6 FP64 instructions / iteration, ONE escape filter per block of 8-16 iterations
                                 and nothing recorded in the loop (needs 8 <= limit <= 2^20 and |c| <= limit^2/2 - 3 limit: an
                                 orbit that fired the filter then provably escapes at once and keeps growing); a stopped lane
                                 stores the position at the start of its last block and a resolve pass re-runs <= 2 block - 1
                                 iterations with the reference's test after every step */
    FRB_KERNEL_TILE = 6       /* image formats with an iteration cap <= 384: one thread per pixel on 8x4 tiles, filter every
                                 iteration, colour in the same kernel (no records, no second kernel); exact for ANY input (it
                                 takes the 7-instruction form where the 6-instruction one is not provably exact) */
};

/* ---- lifecycle ------------------------------------------------------------------------------------------ */

/* Number of usable CUDA devices (0 if none; never fails). */
int frb_device_count(void);
/* Optional eager initialisation. device_mask bit d = use device d; 0 = all visible devices.
 * Creates contexts, streams and work queues so that the first render is not charged for them. */
int frb_init(uint32_t device_mask);
/* The same with an explicit list: a render with device = -1 cuts the frame into n shards and gives shard i to devices[i].
 * A device may appear more than once (its shards are rendered one after the other) — e.g. {0, 0, 1} gives device 0 two thirds
 * of the bands; n <= 32. */
int frb_init_list(const int32_t *devices, uint32_t n);
/* Releases all device and pinned memory, streams and events. Safe to call repeatedly. */
void frb_shutdown(void);
/* Forces the iteration kernel (FRB_KERNEL_*) for subsequent renders of this thread; FRB_KERNEL_AUTO restores
 * the default. Test/bench hook. */
int frb_set_kernel_variant(uint32_t variant);
/* Tuning build only (libfrb_tune.so, -DFRB_TUNING; used by the sweeps under scripts/ and by tests that force rarely taken
 * paths): sets one of the per-frame choices the library otherwise makes itself ("defer_b", "chunk", "fast_shade", ...; "reset"
 * restores the defaults).  The product library reads no environment variable and has no such switch: there this returns
 * FRB_ERR_INVALID. */
int frb_set_tuning(const char *name, int64_t value);

/* ---- render: host-buffer entry points (what the reference's get_image binds to) ---------------------- */

/* Renders the whole frame on device `device` (-1 = spread over all initialised devices by interleaved
 * 16-row bands, one host worker thread per device) and writes width*height*format bytes into the HOST
 * buffer `out` (out_len >= that).  Any host memory works; if `out` is pinned (cudaHostAlloc /
 * cudaHostRegister, 16-byte aligned) the image is written into it over PCIe while the iterations run —
 * by every device into its own bands when device = -1 — instead of being copied afterwards.
 * Replaces get_image's Mandelbrot|Julia arm (src/lib.rs:255-269). */
int frb_render(const FrbConfig *cfg, uint8_t *out, size_t out_len, uint32_t format, int device);

/* Renders one shard of the frame on `device` into a HOST buffer, compact layout (see FrbShard). */
int frb_render_shard(const FrbConfig *cfg, const FrbShard *shard, uint8_t *out, size_t out_len,
                     uint32_t format, int device);

/* Renders one shard of the frame on `device` INTO ITS PLACE in a whole frame: `frame` addresses all width*height*format bytes
 * (frame_len >= that) and only the rows the shard owns are written.  This is the gather of the one-process-per-GPU layout
 * (SURVEY §8e): every process passes the same frame — a frb_shared_frame, or any memory all of them map — and its own shard,
 * and the image is complete when all have returned; no collective, no second copy.  A pinned frame (band_rows a multiple of 16
 * or rows of a multiple of 16 bytes) is written over PCIe while the iterations run.  frb_render(device = -1) does the same
 * with one host thread per device. */
int frb_render_bands(const FrbConfig *cfg, const FrbShard *shard, uint8_t *frame, size_t frame_len,
                     uint32_t format, int device);

/* Parity hook: the raw result of calc::recursive (calc/src/lib.rs:245-257) per pixel of the shard
 * (shard may be NULL = whole frame): iters[p] = returned count (index of the escaping iteration, or
 * `iterations`), zfinal[2p],zfinal[2p+1] = returned position. HOST buffers; zfinal may be NULL. */
int frb_render_counts(const FrbConfig *cfg, const FrbShard *shard, uint32_t *iters, double *zfinal,
                      int device);

/* n independent frames (zoom animation); frame k goes to device k % devices (device = -1) or
 * all to `device`; out holds the frames back to back, each width_k*height_k*format bytes. */
int frb_render_frames(const FrbConfig *cfgs, uint32_t n, uint8_t *out, size_t out_len, uint32_t format,
                      int device);
/* The same for one process of a one-process-per-GPU job: renders only the frames first, first + stride, first + 2 stride, ...
 * (rank and world size) — each into ITS place in `out`, which addresses all n frames (a frb_shared_frame, say), so the animation
 * is complete when every process has returned.  Frames go to the devices in order from a dynamic queue, one host thread per
 * device, two frames in flight on each. */
int frb_render_frames_strided(const FrbConfig *cfgs, uint32_t n, uint8_t *out, size_t out_len, uint32_t format,
                              int device, uint32_t first, uint32_t stride);

/* ---- Barnsley fern: the other arm of get_image (src/lib.rs:271-319, 417-463) --------------------------------- */

/* Replaces get_image's BarnsleyFern arm.  The reference plays the chaos game on T = rayon threads, each with an
 * entropy-seeded generator and its own image (initialised to secondary_color, darkened under every plotted point by
 * subtract_pixel), for iterations / T points each, and reduces the T images with a saturating add; its output
 * therefore depends on the host's thread count and differs from run to run.  Here `threads` is that T (0 = what rayon
 * would use on this host: RAYON_NUM_THREADS or the logical CPUs) and `seed` replaces the entropy: the image is a pure
 * function of (cfg, threads, seed), reproduced bit for bit by oracle/fern_oracle.c.  Each of the T shares is played as
 * ceil(share / 512) independent chains (xoshiro256++ streams seeded seed + (chain+1) * 0xD1342543DE82EF95 via
 * splitmix64; chains after a share's first discard 64 burn-in points).  `shard` (NULL = whole frame) selects the rows
 * written to `out`, compact as for frb_render_shard; every shard plays the whole game (10^7 points are microseconds).
 * frb_render* with algo = FRB_ALGO_BARNSLEY_FERN keep the per-pixel semantics (get_recursive_pixel returns black). */
int frb_render_fern(const FrbConfig *cfg, uint32_t threads, uint64_t seed, const FrbShard *shard, uint8_t *out,
                    size_t out_len, uint32_t format, int device);
/* rayon::current_num_threads() as the reference would see it on this host. Host-only. */
uint32_t frb_fern_default_threads(void);

/* ---- render: device-buffer entry points (no host copies; bench `value`, GPU-resident consumers) ----- */

/* Same as frb_render_shard but `dev_out` is DEVICE memory on `device` (>= rows*width*format bytes,
 * 16-byte aligned). `stream` is a cudaStream_t (NULL = the library's own stream for that device).
 * Returns after enqueueing; results are complete when `stream` has drained. Stats (iterations etc.)
 * are harvested by frb_sync_stats().  Renders on ONE stream run strictly in order; renders issued on two
 * different streams use separate sets of the library's device buffers and may overlap on the device. */
int frb_render_device_async(const FrbConfig *cfg, const FrbShard *shard, void *dev_out, size_t out_len,
                            uint32_t format, int device, void *stream);
/* Waits for the last *_async render of this thread and fills its statistics. */
int frb_sync_stats(FrbStats *stats);
/* Same for the *_async render issued `back` calls before the last one by this thread (0 = the last; the
 * library keeps the 32 most recent), so a caller can keep the next frame enqueued while harvesting. */
int frb_async_stats(uint32_t back, FrbStats *stats);

/* ---- render without allocating (no counterpart in the reference) ------------------------------------------------- */

/* No counterpart in the reference (its get_image allocates a fresh Vec per call, src/lib.rs:266): an addition for hosts that
 * must not, or cannot, allocate GPU memory at the moment they want a frame.  Every other render entry point grows the
 * library's device buffers with cudaMalloc when a frame is larger than the last; that is out of the question for a plug-in
 * inside a process whose own pool allocator has already taken all of HBM (a cudaMalloc there fails, or starves the host's
 * next allocation), for an interactive loop that cannot afford the device-wide synchronisation cudaMalloc / cudaFree imply
 * each time the window is resized, and for code running where the CUDA allocator must not be entered (inside another
 * library's callback, under a lock its allocator also takes).
 *
 * This call renders the WHOLE frame of `cfg` into the caller's DEVICE buffer `dev_out` (on `device`, 16-byte aligned,
 * out_len >= width*height*format) on `stream` (a cudaStream_t; NULL = the legacy default stream) and returns after
 * enqueueing three kernels; the image is complete when `stream` has drained.  It calls no cudaMalloc, cudaFree,
 * cudaHostAlloc, stream / event creation or other allocating routine, on any call.  Its scratch memory (coordinate tables,
 * per-pixel count and final-z records, work counter, statistics words) is ONE file-scope `static __device__` array in
 * csrc/frb_noalloc.cu, so:
 *   - LIMIT: width * height <= 1 048 576 pixels and width + height <= 16 384; anything larger is refused with
 *     FRB_ERR_TOO_LARGE and a message naming the limit (use frb_render_device_async there).  The array is 20.1 MiB; every GPU
 *     that loads this library's code holds it until the process exits, used or not;
 *   - the FIRST call on each device does the set-up for that device (looks up the array's device address with
 *     cudaGetSymbolAddress while that device is current, queries occupancy, loads the kernels' code).  Make that first call
 *     early — in particular before any stream capture is begun;
 *   - all calls on one device share the array, and it keeps the previous call's contents (the first kernel of each call
 *     clears the counter and statistics words, every other word is written before it is read within the call).  Two calls
 *     whose kernels overlap on the GPU — different streams, or two host threads on unordered streams — RACE on it and both
 *     images are garbage.  Order them: same stream, or an event between them.  Other entry points use other memory and
 *     may run beside it.
 * Pixels are those of frb_render_device_async, byte for byte (it runs the 7-instruction persistent kernel, exact for any
 * input, then the same shade kernel); no statistics are harvested (frb_sync_stats does not see this call). */
int frb_render_static_async(const FrbConfig *cfg, void *dev_out, size_t out_len, uint32_t format, int device, void *stream);

/* ---- headless frame server (the reference GUI's render thread without the window) -------- */

/* Mirrors src/gui.rs:37-48 (request_redraw, latest-wins coalescing through working/try_redraw), :56-82 (render thread:
 * config -> get_image -> RGBA8 frame -> shared slot) and :110-117 (the UI takes the frame and re-requests).  The GL
 * texture upload is left out (no graphics stack); the caller receives the RGBA frame. */
typedef struct FrbServer FrbServer;
/* Renderer hook: fill `rgba` (width*height*4 bytes) for `cfg`, return FRB_OK or an error.  NULL = frb_render(RGBA32). */
typedef int (*FrbServerRenderFn)(const FrbConfig *cfg, uint8_t *rgba, size_t len, void *user);
typedef struct FrbServerStats {
    uint64_t requests;        /* frb_server_request calls + re-requests issued on poll */
    uint64_t coalesced;       /* requests absorbed because a render was in flight (try_redraw) */
    uint64_t frames_rendered;
    uint32_t working;
    uint32_t try_redraw;
} FrbServerStats;

FrbServer *frb_server_create(int device, FrbServerRenderFn renderer, void *user);
void frb_server_destroy(FrbServer *server);
/* The state changed (GUI: `config != previous_state`): render it, or, if a render is in flight, remember that the
 * newest state must be rendered once the current frame has been taken.  Never blocks. */
int frb_server_request(FrbServer *server, const FrbConfig *cfg);
/* Takes the finished frame if there is one: returns 1 and copies width*height*4 bytes, 0 if none, < 0 if the last render
 * failed (or `len` is too small).  rgba == NULL peeks: reports the size and leaves the frame in the slot.
 * timeout_ms: 0 = do not wait, < 0 = wait for a frame. */
int frb_server_poll(FrbServer *server, uint8_t *rgba, size_t len, uint32_t *width, uint32_t *height, int timeout_ms);
/* Like frb_server_poll, without the copy: hands out the finished frame itself — `*rgba` points at width*height*4 bytes
 * of the server's (pinned) memory and stays valid until the next frb_server_acquire / frb_server_poll that returns a
 * frame, or frb_server_destroy.  This is the reference's `image.lock().unwrap().take()` (src/gui.rs:110): ownership of
 * the frame moves to the UI side; the render thread keeps working in its other two buffers.  Returns 1 / 0 / <0 as poll. */
int frb_server_acquire(FrbServer *server, const uint8_t **rgba, uint32_t *width, uint32_t *height, int timeout_ms);
int frb_server_stats(FrbServer *server, FrbServerStats *stats);
const char *frb_server_last_error(FrbServer *server);

/* ---- pinned host memory (optional; lets a binding use the fast output path without linking CUDA itself) ---------- */

/* Page-locked host memory, usable from every device: a destination allocated here is written by the GPU while the
 * iterations run (frb_render / frb_render_shard) instead of being copied afterwards.  NULL on failure (see
 * frb_last_error).  Release with frb_host_free.  On the reference side this would back the Vec<RGB> that get_image
 * returns (src/lib.rs:253) or the GUI's frame (src/gui.rs:62-75) through a custom allocator / Box<[u8]>. */
void *frb_host_alloc(size_t bytes);
void frb_host_free(void *p);
/* Pins caller-owned memory in place (e.g. a long-lived frame buffer) / undoes it.  Pinning costs about a millisecond
 * per 10 MB: worth it for buffers that are rendered into many times. */
int frb_host_register(void *p, size_t bytes);
int frb_host_unregister(void *p);

/* ---- shared frame: one host image written by several processes (one per GPU) ------------------------------------------ */

/* A POSIX shared-memory image that every rank of a one-process-per-GPU job maps and pins (cudaHostRegister), so that each
 * rank's frb_render_bands lands in the same host frame — the reference's single collected Vec<RGB> (src/lib.rs:266-267) —
 * without a gather step.  `name` identifies the frame on this host ("/frb-<job>"), `bytes` its size, `ranks` the number of
 * processes; rank 0 creates it, the others attach (waiting up to ~30 s for it to appear).  Works without a CUDA device (the
 * mapping is then simply not pinned), which is how the host logic is tested on CPU.  NULL on failure (frb_last_error). */
typedef struct FrbSharedFrame FrbSharedFrame;
FrbSharedFrame *frb_shared_frame_open(const char *name, size_t bytes, uint32_t ranks, uint32_t rank);
/* The image bytes (page aligned). */
uint8_t *frb_shared_frame_data(FrbSharedFrame *frame);
/* 1 if the mapping is page-locked (a GPU can write it directly). */
int frb_shared_frame_pinned(FrbSharedFrame *frame);
/* Returns when all `ranks` processes have called it the same number of times (sense-reversing counter in the shared header;
 * spins, then yields): "the whole frame is on the host".  FRB_ERR_INVALID after ~60 s without the others, or at once when this
 * rank's mapping turns out to be a removed or replaced file (a crashed job's left-over under the same name).  It is ONE phase:
 * a rank that returns may start the next frame's writes at once, so a consumer that reads frame k in place while the ranks go
 * on needs a second barrier call after its read (or double-buffers with two frames). */
int frb_shared_frame_barrier(FrbSharedFrame *frame);
/* Unpins and unmaps; rank 0 also removes the name. */
void frb_shared_frame_close(FrbSharedFrame *frame);

/* ---- diagnostics ---------------------------------------------------------------------------------------- */

/* Message for the last error on this thread ("" if none). Valid until the next failing call on this thread. */
const char *frb_last_error(void);
/* Statistics of the last blocking render on this thread. */
int frb_last_stats(FrbStats *stats);
/* ABI version of the loaded library (== FRB_ABI_VERSION). */
uint32_t frb_abi_version(void);

/* Measures the FP64 pipe of `device`: dependent-chain-free DFMA / DMUL / DADD streams on every SM.
 * Results in instructions per second (thread-level, i.e. lanes x warp-instructions). This is the
 * roofline denominator for this path (MEASURED_PEAKS.json has no FP64 entry). */
typedef struct FrbFp64Peak {
    double dfma_per_s;
    double dmul_per_s;
    double dadd_per_s;
    double mixed6_per_s;  /* the 6-instruction iteration body as an independent stream (no exits) */
    double sm_clock_mhz;  /* SM clock implied by the DFMA rate: dfma_per_s / (sm_count x 64 FP64 lanes) */
    uint32_t sm_count;
    uint32_t reserved;
} FrbFp64Peak;
int frb_measure_fp64_peak(int device, FrbFp64Peak *peak);

#ifdef __cplusplus
}
#endif
#endif /* FRB_H_ */
