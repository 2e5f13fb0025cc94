// build.rs -- builds libdis_engine.so (this repository's csrc/) with nvcc for sm_100a and links it.
// UNVERIFIED: written without a Rust toolchain; see ../README.md.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let src = PathBuf::from(env::var("DIS_B200_SRC").expect("DIS_B200_SRC = path to the B200 engine repository"));
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let csrc = src.join("distributed_inference_server_b200/csrc");
    // the same list as distributed_inference_server_b200/build.py
    let files = [
        "k_elementwise.cu", "k_gemv.cu", "k_attn_paged.cu", "k_attn_prefill_tc.cu", "k_gemm_tcgen05.cu", "k_gemm_persist.cu", "k_tp.cu", "k_tp_sample.cu",
        "engine.cu", "api.cpp", "tp.cpp", "worker.cpp", "serving.cpp", "prefix.cpp",
    ];
    let mut cmd = Command::new(env::var("NVCC").unwrap_or_else(|_| "nvcc".into()));
    cmd.args([
        "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr",
        "-shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-x", "cu",
    ]);
    for f in files {
        cmd.arg(csrc.join(f));
    }
    cmd.arg("-o").arg(out.join("libdis_engine.so")).arg("-ldl");
    let status = cmd.status().expect("failed to run nvcc");
    assert!(status.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=dis_engine");
    println!("cargo:rerun-if-changed={}", csrc.display());
    println!("cargo:rerun-if-changed={}", src.join("include/dis_engine.h").display());
}
