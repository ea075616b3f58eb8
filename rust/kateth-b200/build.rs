// Builds libkateth_b200.so (hand-written sm_100a CUDA + the host pairing) with nvcc and links it.
// The sources live two directories up (kateth_b200/csrc); KATETH_B200_LIB_DIR skips the build and links a prebuilt library.
use std::{env, path::PathBuf, process::Command};

fn main() {
    println!("cargo:rerun-if-env-changed=KATETH_B200_LIB_DIR");
    if let Ok(dir) = env::var("KATETH_B200_LIB_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-lib=dylib=kateth_b200");
        return;
    }
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    let src = root.join("kateth_b200/csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let obj = out.join("host_bls.o");
    let ok = Command::new("g++")
        .args(["-O3", "-march=x86-64-v3", "-fPIC", "-std=c++17", "-c"])
        .arg(src.join("host_bls.cpp")).arg("-o").arg(&obj)
        .status().expect("g++ not found").success();
    assert!(ok, "host_bls.cpp failed to compile");
    let ok = Command::new(env::var("NVCC").unwrap_or_else(|_| "nvcc".into()))
        .args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-shared"])
        .arg(src.join("engine.cu")).arg(&obj)
        .arg("-o").arg(out.join("libkateth_b200.so")).arg("-lcudart")
        .status().expect("nvcc not found").success();
    assert!(ok, "engine.cu failed to compile for sm_100a");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=kateth_b200");
    println!("cargo:rerun-if-changed={}", src.display());
    println!("cargo:rerun-if-changed={}", root.join("include/kateth_b200.h").display());
}
