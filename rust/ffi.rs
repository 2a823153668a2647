//! ffi.rs — the binding a maintainer adds to the reference crate (`src/ffi.rs`) to route get_image's
//! Mandelbrot | Julia arm (src/lib.rs:255-269) through libfrb.so.  Mirrors include/frb.h 1:1.
//! NOT compiled here: this image has no rustc/cargo.  The same ABI is exercised from C++ and Python.
use fractal_renderer_calc::{Algo, Config, RGB};
use std::ffi::CStr;
use std::os::raw::{c_char, c_int};

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

#[link(name = "frb")]
extern "C" {
    pub fn frb_init(device_mask: u32) -> c_int;
    pub fn frb_shutdown();
    pub fn frb_device_count() -> c_int;
    pub fn frb_render(cfg: *const FrbConfig, out: *mut u8, out_len: usize, format: u32, device: c_int) -> c_int;
    pub fn frb_render_fern(cfg: *const FrbConfig, threads: u32, seed: u64, shard: *const core::ffi::c_void, out: *mut u8,
                           out_len: usize, format: u32, device: c_int) -> c_int;
    pub fn frb_last_error() -> *const c_char;
    pub fn frb_host_alloc(bytes: usize) -> *mut core::ffi::c_void;
    pub fn frb_host_free(p: *mut core::ffi::c_void);
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
    let rc = unsafe { frb_render_fern(&cfg, threads, seed, core::ptr::null(), image.as_mut_ptr() as *mut u8, n * 3, FRB_RGB24, 0) };
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
                frb_host_free(self.ptr as *mut core::ffi::c_void);
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
            unsafe { frb_host_free(self.ptr as *mut core::ffi::c_void) };
        }
    }
}
