// build.rs -- compiles the CUDA sources of ../whisper_burn_b200/csrc for sm_100a with nvcc and links
// the resulting static-cudart shared library.  No Triton, no wgpu/CubeCL, no CPU fallback.
use std::env;
use std::path::PathBuf;
use std::process::Command;

fn main() {
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let csrc = manifest.join("../whisper_burn_b200/csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    let lib = out.join("libwq4.so");
    let sources = ["wq4_api.cu", "wq4_tensor.cu", "wq4_gemv.cu", "wq4_skinny.cu", "wq4_gemm.cu", "wq4_gemm2.cu", "wq4_model_kernels.cu", "wq4_decode.cu", "wq4_attn.cu", "wq4_mel.cu", "wq4_whisper.cu"];
    let mut cmd = Command::new(&nvcc);
    cmd.args([
        "-arch=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-shared", "-cudart", "static",
        "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC,-fvisibility=hidden", "-o",
    ]);
    cmd.arg(&lib);
    for s in sources {
        let p = csrc.join(s);
        println!("cargo:rerun-if-changed={}", p.display());
        cmd.arg(p);
    }
    println!("cargo:rerun-if-changed={}", manifest.join("../include/wq4.h").display());
    let status = cmd.status().expect("failed to run nvcc (set NVCC=/path/to/nvcc)");
    assert!(status.success(), "nvcc -arch=sm_100a failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=wq4");
}
