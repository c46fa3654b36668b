// build.rs — compiles the hand-written kernels (csrc/rtc_api.cu includes every .cuh) for sm_100a and links them
// statically. NOT COMPILED IN THE BUILD IMAGE OF THIS REPOSITORY (no rustc/cargo there): shipped as source.
//
//   RTC_CUDA_ROOT     directory holding `csrc/` and `include/` (default: next to this crate — copy
//                     raytracer-challenge-rs_b200/csrc and include/ of the backend repository there)
//   NVCC              nvcc binary (default: nvcc on PATH)
//   CUDA_HOME         for libcudart (default /usr/local/cuda)
//   --features prebuilt + RTC_CUDA_LIB_DIR   link an existing librtc_cuda.so instead
use std::{env, path::PathBuf, process::Command};

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let cuda_home = env::var("CUDA_HOME").unwrap_or_else(|_| "/usr/local/cuda".into());
    println!("cargo:rustc-link-search=native={}/lib64", cuda_home);
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=stdc++");
    if env::var("CARGO_FEATURE_PREBUILT").is_ok() {
        let dir = env::var("RTC_CUDA_LIB_DIR").expect("feature `prebuilt` needs RTC_CUDA_LIB_DIR");
        println!("cargo:rustc-link-search=native={}", dir);
        println!("cargo:rustc-link-lib=dylib=rtc_cuda");
        println!("cargo:rerun-if-env-changed=RTC_CUDA_LIB_DIR");
        return;
    }
    let root = env::var("RTC_CUDA_ROOT")
        .map(PathBuf::from)
        .unwrap_or_else(|_| PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()));
    let src = root.join("csrc").join("rtc_api.cu");
    let obj = out.join("rtc_api.o");
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "nvcc".into());
    // --fmad=false / -ffp-contract=off: Rust never fuses a*b+c, and the frames must be the CPU render's bit for bit
    let status = Command::new(&nvcc)
        .args(&["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "--fmad=false", "-std=c++17"])
        .args(&["-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden"])
        .arg("-I")
        .arg(root.join("include"))
        .arg("-c")
        .arg(&src)
        .arg("-o")
        .arg(&obj)
        .status()
        .expect("failed to run nvcc");
    assert!(status.success(), "nvcc failed on {}", src.display());
    let lib = out.join("librtc_cuda.a");
    let status = Command::new("ar").arg("crs").arg(&lib).arg(&obj).status().expect("failed to run ar");
    assert!(status.success());
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=static=rtc_cuda");
    println!("cargo:rerun-if-changed={}", root.join("csrc").display());
    println!("cargo:rerun-if-changed={}", root.join("include").join("rtc.h").display());
    println!("cargo:rerun-if-env-changed=RTC_CUDA_ROOT");
}
