// build.rs — replaces the reference's dead `gpu` block (build.rs:4-52): compiles the CUDA library with nvcc for
// sm_100a and links it.  NOT run here (no cargo on this image); `fractal_renderer_b200/csrc/Makefile` does the same, and
// tests/test_rust_shim.py checks that SOURCES and the flags below agree with that Makefile.
use std::{env, path::PathBuf, process::Command};

/// The translation units of libfrb.so == `SRC` in fractal_renderer_b200/csrc/Makefile.
const SOURCES: &[&str] = &["frb_api.cu", "frb_peak.cu", "frb_server.cu", "frb_shared.cu", "frb_noalloc.cu"];

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let src = PathBuf::from(env::var("FRB_SRC").unwrap_or_else(|_| "fractal_renderer_b200/csrc".into()));
    let lib = out.join("libfrb.so");
    let status = Command::new(env::var("NVCC").unwrap_or_else(|_| "nvcc".into()))
        .args(["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false"])
        .args(["-Xcompiler", "-fPIC", "-cudart", "static", "-shared", "-o"])
        .arg(&lib)
        .args(SOURCES.iter().map(|f| src.join(f)))
        .status()
        .expect("nvcc not found");
    assert!(status.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=frb");
    println!("cargo:rerun-if-changed={}", src.display());
}
