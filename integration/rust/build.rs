// build.rs — compiles the CUDA backend (rust-ray-tracing_b200/csrc, vendored under rtb200/ in the reference's crate root) with
// nvcc for sm_100a and links it.  Replaces nothing in the reference: the crate has no build script (Cargo.toml:1-22).
//
// Layout expected at the crate root:   rtb200/include/rtb200.h   rtb200/csrc/{rtb200.cu, rt_kernels.cuh, rt_device.cuh, rt_filter.h}
// `-fmad=false` is required: the kernels place every fused multiply-add explicitly where the reference calls mul_add
// (src/geometry.rs:435,467; src/objects.rs:257).  The Rust toolchain is absent on the box this backend was built and tested
// on; the same nvcc command line is what rust-ray-tracing_b200/csrc/Makefile runs there.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let src = PathBuf::from("rtb200/csrc/rtb200.cu");
    let lib = out.join("librtb200.so");
    let status = Command::new(env::var("NVCC").unwrap_or_else(|_| "nvcc".into()))
        .args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
               "-Xcompiler", "-fPIC", "-shared", "-o"])
        .arg(&lib)
        .arg(&src)
        .status()
        .expect("nvcc not found (set NVCC)");
    assert!(status.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=rtb200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", out.display());
    println!("cargo:rerun-if-changed=rtb200/csrc");
    println!("cargo:rerun-if-changed=rtb200/include/rtb200.h");
}
