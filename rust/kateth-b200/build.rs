// Builds libkateth_b200.so (hand-written sm_100a CUDA + the host pairing + the NCCL group front end) and links it.
// The sources live two directories up (kateth_b200/csrc) and their Makefile is the single recipe: engine.cu and msm_ba.cu with
// `nvcc -gencode arch=compute_100a,code=sm_100a`, host_bls.cpp and group.cpp with g++.  KATETH_B200_LIB_DIR skips the build and
// links a prebuilt library.
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
    let ok = Command::new("make").arg("-C").arg(&src).arg("-j4").arg("-s").status().expect("make not found").success();
    assert!(ok, "kateth_b200/csrc failed to build for sm_100a (needs nvcc 12.8+ and g++)");
    println!("cargo:rustc-link-search=native={}", root.join("kateth_b200").display());
    println!("cargo:rustc-link-lib=dylib=kateth_b200");
    println!("cargo:rerun-if-changed={}", src.display());
    println!("cargo:rerun-if-changed={}", root.join("include/kateth_b200.h").display());
}
