//! ffi.rs — the binding a maintainer adds to the reference crate (`src/ffi.rs`) to route get_image's
//! Mandelbrot | Julia arm (src/lib.rs:255-269) through libfrb.so.  Mirrors include/frb.h 1:1 (every struct and every function;
//! tests/test_rust_shim.py holds the two files against each other).
//! NOT compiled here: this image has no rustc/cargo.  The same ABI is exercised from C++ and Python.
use fractal_renderer_calc::{Algo, Config, RGB};
use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)]
#[derive(Clone, Copy)]
pub struct FrbImaginary {
    pub re: f64,
    pub im: f64,
}
#[repr(C)]
#[derive(Clone, Copy)]
pub struct FrbRGB {
    pub r: u8,
    pub g: u8,
    pub b: u8,
}
/// calc::Config field for field (calc/src/lib.rs:21-37); 104 bytes.
#[repr(C)]
pub struct FrbConfig {
    pub algo: u32,
    pub width: u32,
    pub height: u32,
    pub iterations: u32,
    pub limit: f64,
    pub stable_limit: f64,
    pub pos: FrbImaginary,
    pub scale: FrbImaginary,
    pub exposure: f64,
    pub inside: u8,
    pub smooth: u8,
    pub primary_color: FrbRGB,
    pub secondary_color: FrbRGB,
    pub color_weight: f64,
    pub julia_set: FrbImaginary,
}
pub const FRB_RGB24: u32 = 3;
pub const FRB_RGBA32: u32 = 4;

/// Interleaved row bands of one frame (include/frb.h: FrbShard); {0, 1, 0, 0} = the whole frame.
#[repr(C)]
#[derive(Clone, Copy)]
pub struct FrbShard {
    pub index: u32,
    pub count: u32,
    pub band_rows: u32,
    pub reserved: u32,
}
/// Statistics of a render (include/frb.h: FrbStats); 56 bytes.
#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct FrbStats {
    pub pixels: u64,
    pub iterations: u64,
    pub lane_iterations: u64,
    pub iterate_ms: f32,
    pub shade_ms: f32,
    pub device_ms: f32,
    pub total_ms: f32,
    pub kernel_launches: u32,
    pub devices: u32,
    pub kernel_variant: u32,
    pub reserved: u32,
}
#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct FrbServerStats {
    pub requests: u64,
    pub coalesced: u64,
    pub frames_rendered: u64,
    pub working: u32,
    pub try_redraw: u32,
}
#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct FrbFp64Peak {
    pub dfma_per_s: f64,
    pub dmul_per_s: f64,
    pub dadd_per_s: f64,
    pub mixed6_per_s: f64,
    pub sm_clock_mhz: f64,
    pub sm_count: u32,
    pub reserved: u32,
}
/// Opaque handles (the library owns them).
#[repr(C)]
pub struct FrbServer {
    _private: [u8; 0],
}
#[repr(C)]
pub struct FrbSharedFrame {
    _private: [u8; 0],
}
/// Renderer hook of the frame server; None = frb_render(RGBA32).
pub type FrbServerRenderFn = Option<unsafe extern "C" fn(cfg: *const FrbConfig, rgba: *mut u8, len: usize, user: *mut c_void) -> c_int>;

// Every function include/frb.h declares, in the header's order (tests/test_rust_shim.py compares the two files).
#[link(name = "frb")]
extern "C" {
    pub fn frb_config_default(cfg: *mut FrbConfig, algo: u32);
    pub fn frb_rgb_new(r: u8, b: u8, g: u8) -> FrbRGB;
    pub fn frb_shard_rows(height: u32, shard: *const FrbShard) -> u32;
    pub fn frb_shard_row_to_y(height: u32, shard: *const FrbShard, local_row: u32) -> u32;
    pub fn frb_device_count() -> c_int;
    pub fn frb_init(device_mask: u32) -> c_int;
    pub fn frb_init_list(devices: *const i32, n: u32) -> c_int;
    pub fn frb_shutdown();
    pub fn frb_set_kernel_variant(variant: u32) -> c_int;
    pub fn frb_set_tuning(name: *const c_char, value: i64) -> c_int;
    pub fn frb_render(cfg: *const FrbConfig, out: *mut u8, out_len: usize, format: u32, device: c_int) -> c_int;
    pub fn frb_render_shard(cfg: *const FrbConfig, shard: *const FrbShard, out: *mut u8, out_len: usize, format: u32, device: c_int) -> c_int;
    pub fn frb_render_bands(cfg: *const FrbConfig, shard: *const FrbShard, frame: *mut u8, frame_len: usize, format: u32, device: c_int) -> c_int;
    pub fn frb_render_counts(cfg: *const FrbConfig, shard: *const FrbShard, iters: *mut u32, zfinal: *mut f64, device: c_int) -> c_int;
    pub fn frb_render_frames(cfgs: *const FrbConfig, n: u32, out: *mut u8, out_len: usize, format: u32, device: c_int) -> c_int;
    pub fn frb_render_frames_strided(cfgs: *const FrbConfig, n: u32, out: *mut u8, out_len: usize, format: u32, device: c_int, first: u32,
                                     stride: u32) -> c_int;
    pub fn frb_render_fern(cfg: *const FrbConfig, threads: u32, seed: u64, shard: *const FrbShard, out: *mut u8, out_len: usize, format: u32,
                           device: c_int) -> c_int;
    pub fn frb_fern_default_threads() -> u32;
    pub fn frb_render_device_async(cfg: *const FrbConfig, shard: *const FrbShard, dev_out: *mut c_void, out_len: usize, format: u32,
                                   device: c_int, stream: *mut c_void) -> c_int;
    pub fn frb_sync_stats(stats: *mut FrbStats) -> c_int;
    pub fn frb_async_stats(back: u32, stats: *mut FrbStats) -> c_int;
    pub fn frb_render_static_async(cfg: *const FrbConfig, dev_out: *mut c_void, out_len: usize, format: u32, device: c_int,
                                   stream: *mut c_void) -> c_int;
    pub fn frb_server_create(device: c_int, renderer: FrbServerRenderFn, user: *mut c_void) -> *mut FrbServer;
    pub fn frb_server_destroy(server: *mut FrbServer);
    pub fn frb_server_request(server: *mut FrbServer, cfg: *const FrbConfig) -> c_int;
    pub fn frb_server_poll(server: *mut FrbServer, rgba: *mut u8, len: usize, width: *mut u32, height: *mut u32, timeout_ms: c_int) -> c_int;
    pub fn frb_server_acquire(server: *mut FrbServer, rgba: *mut *const u8, width: *mut u32, height: *mut u32, timeout_ms: c_int) -> c_int;
    pub fn frb_server_stats(server: *mut FrbServer, stats: *mut FrbServerStats) -> c_int;
    pub fn frb_server_last_error(server: *mut FrbServer) -> *const c_char;
    pub fn frb_host_alloc(bytes: usize) -> *mut c_void;
    pub fn frb_host_free(p: *mut c_void);
    pub fn frb_host_register(p: *mut c_void, bytes: usize) -> c_int;
    pub fn frb_host_unregister(p: *mut c_void) -> c_int;
    pub fn frb_shared_frame_open(name: *const c_char, bytes: usize, ranks: u32, rank: u32) -> *mut FrbSharedFrame;
    pub fn frb_shared_frame_data(frame: *mut FrbSharedFrame) -> *mut u8;
    pub fn frb_shared_frame_pinned(frame: *mut FrbSharedFrame) -> c_int;
    pub fn frb_shared_frame_barrier(frame: *mut FrbSharedFrame) -> c_int;
    pub fn frb_shared_frame_close(frame: *mut FrbSharedFrame);
    pub fn frb_last_error() -> *const c_char;
    pub fn frb_last_stats(stats: *mut FrbStats) -> c_int;
    pub fn frb_abi_version() -> u32;
    pub fn frb_measure_fp64_peak(device: c_int, peak: *mut FrbFp64Peak) -> c_int;
}

impl From<&Config> for FrbConfig {
    fn from(c: &Config) -> Self {
        let rgb = |v: RGB| FrbRGB { r: v.r, g: v.g, b: v.b }; // STORED fields; RGB::new's swap already happened
        FrbConfig {
            algo: match c.algo {
                Algo::Mandelbrot => 0,
                Algo::BarnsleyFern => 1,
                Algo::Julia => 2,
            },
            width: c.width,
            height: c.height,
            iterations: c.iterations,
            limit: c.limit,
            stable_limit: c.stable_limit,
            pos: FrbImaginary { re: c.pos.re, im: c.pos.im },
            scale: FrbImaginary { re: c.scale.re, im: c.scale.im },
            exposure: c.exposure,
            inside: c.inside as u8,
            smooth: c.smooth as u8,
            primary_color: rgb(c.primary_color),
            secondary_color: rgb(c.secondary_color),
            color_weight: c.color_weight,
            julia_set: FrbImaginary { re: c.julia_set.re, im: c.julia_set.im },
        }
    }
}

/// Drop-in body of get_image's `Algo::Mandelbrot | Algo::Julia` arm.  `RGB` must be `#[repr(C)]`
/// (3 packed bytes r,g,b — which src/lib.rs:13-15 and src/gui.rs:62-64 already assume).
pub fn get_image_b200(config: &Config) -> Vec<RGB> {
    let n = config.width as usize * config.height as usize;
    let mut image: Vec<RGB> = Vec::with_capacity(n);
    let cfg = FrbConfig::from(config);
    // device -1: interleaved 16-row bands over every visible B200
    let rc = unsafe { frb_render(&cfg, image.as_mut_ptr() as *mut u8, n * 3, FRB_RGB24, -1) };
    if rc != 0 {
        let msg = unsafe { CStr::from_ptr(frb_last_error()) }.to_string_lossy().into_owned();
        panic!("frb_render failed ({}): {}", rc, msg); // the reference's error style is expect()/unwrap()
    }
    unsafe { image.set_len(n) };
    image
}

/// Drop-in body of get_image's `Algo::BarnsleyFern` arm (src/lib.rs:271-319): T = rayon's thread count as before, the
/// entropy the reference feeds SmallRng (src/lib.rs:428) becomes the seed of the device-side generators.
pub fn get_image_fern_b200(config: &Config) -> Vec<RGB> {
    use rand::RngCore;
    let n = config.width as usize * config.height as usize;
    let mut image: Vec<RGB> = Vec::with_capacity(n);
    let cfg = FrbConfig::from(config);
    let threads = rayon::current_num_threads() as u32;
    let seed = rand::rngs::OsRng.next_u64();
    let rc = unsafe { frb_render_fern(&cfg, threads, seed, core::ptr::null::<FrbShard>(), image.as_mut_ptr() as *mut u8, n * 3, FRB_RGB24, 0) };
    if rc != 0 {
        let msg = unsafe { CStr::from_ptr(frb_last_error()) }.to_string_lossy().into_owned();
        panic!("frb_render_fern failed ({}): {}", rc, msg);
    }
    unsafe { image.set_len(n) };
    image
}

/// A page-locked RGBA frame for callers that render many frames into the same memory — the GUI's render thread
/// (src/gui.rs:56-82): the GPU writes a pinned destination while it iterates, so nothing is copied after the frame.
/// `render` replaces `get_image(&config)` + the RGB->RGBA expansion of src/gui.rs:62-75; `as_rgba` is what
/// `ColorImage::from_rgba_unmultiplied` takes.
pub struct PinnedFrame {
    ptr: *mut u8,
    cap: usize,
    len: usize,
}

unsafe impl Send for PinnedFrame {}

impl PinnedFrame {
    pub fn new() -> Self {
        PinnedFrame { ptr: core::ptr::null_mut(), cap: 0, len: 0 }
    }
    fn reserve(&mut self, bytes: usize) {
        if bytes <= self.cap {
            return;
        }
        unsafe {
            if !self.ptr.is_null() {
                frb_host_free(self.ptr as *mut c_void);
            }
            let want = bytes + bytes / 8 + 4096; // head-room: window resizes should not re-pin every frame
            self.ptr = frb_host_alloc(want) as *mut u8;
            if self.ptr.is_null() {
                let msg = CStr::from_ptr(frb_last_error()).to_string_lossy().into_owned();
                panic!("frb_host_alloc failed: {}", msg);
            }
            self.cap = want;
        }
    }
    pub fn render(&mut self, config: &Config) {
        let n = config.width as usize * config.height as usize * 4;
        self.reserve(n);
        let cfg = FrbConfig::from(config);
        let rc = unsafe { frb_render(&cfg, self.ptr, n, FRB_RGBA32, -1) };
        if rc != 0 {
            let msg = unsafe { CStr::from_ptr(frb_last_error()) }.to_string_lossy().into_owned();
            panic!("frb_render failed ({}): {}", rc, msg);
        }
        self.len = n;
    }
    pub fn as_rgba(&self) -> &[u8] {
        if self.len == 0 {
            return &[];
        }
        unsafe { core::slice::from_raw_parts(self.ptr, self.len) }
    }
}

impl Drop for PinnedFrame {
    fn drop(&mut self) {
        if !self.ptr.is_null() {
            unsafe { frb_host_free(self.ptr as *mut c_void) };
        }
    }
}

/// The GUI's render thread without its thread (src/gui.rs:37-48, 56-82, 110-117): `request` is request_redraw / try_redraw
/// (latest state wins while a frame is in flight), `take` is `image.lock().unwrap().take()` — the finished RGBA frame, still
/// in the server's pinned memory, valid until the next `take` that returns a frame.
pub struct FrameServer {
    raw: *mut FrbServer,
}

unsafe impl Send for FrameServer {}

impl FrameServer {
    pub fn new() -> Self {
        let raw = unsafe { frb_server_create(-1, None, core::ptr::null_mut()) };
        if raw.is_null() {
            let msg = unsafe { CStr::from_ptr(frb_last_error()) }.to_string_lossy().into_owned();
            panic!("frb_server_create failed: {}", msg);
        }
        FrameServer { raw }
    }
    pub fn request(&self, config: &Config) {
        let cfg = FrbConfig::from(config);
        let rc = unsafe { frb_server_request(self.raw, &cfg) };
        assert!(rc == 0, "frb_server_request failed ({})", rc);
    }
    /// (rgba, width, height) of the newest finished frame, if there is one.
    pub fn take(&self) -> Option<(&[u8], u32, u32)> {
        let mut p: *const u8 = core::ptr::null();
        let (mut w, mut h) = (0u32, 0u32);
        let rc = unsafe { frb_server_acquire(self.raw, &mut p, &mut w, &mut h, 0) };
        if rc < 0 {
            let msg = unsafe { CStr::from_ptr(frb_server_last_error(self.raw)) }.to_string_lossy().into_owned();
            panic!("render failed ({}): {}", rc, msg);
        }
        if rc == 0 {
            return None;
        }
        Some((unsafe { core::slice::from_raw_parts(p, w as usize * h as usize * 4) }, w, h))
    }
}

impl Drop for FrameServer {
    fn drop(&mut self) {
        unsafe { frb_server_destroy(self.raw) };
    }
}
