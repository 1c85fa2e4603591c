// build.rs — what quantization/build.rs becomes with the `cuda` cargo feature.  SYNTAX-REVIEWED ONLY: there is no
// Rust toolchain in the image this repository is built in; the same C-ABI is exercised through ctypes by tests/.
//
// The reference's build script (quantization/build.rs:18-40) hands cpp/sse.c, cpp/avx2.c or cpp/neon.c to the `cc`
// crate and links the static library `simd_utils`.  With the feature on, none of those are built (the GPU path has
// no CPU fallback) and one of two things happens:
//
//   QZ_CUDA_DIR set   -> link the prebuilt libqz_cuda.so found there (python -m quantization_b200.build makes it)
//   otherwise         -> compile quantization_b200/csrc/*.cu with nvcc for sm_100a, exactly as build.py does
//                        (QZ_CUDA_SRC = the directory holding csrc/ and ../include, NVCC = the compiler to use)
use std::{env, path::PathBuf, process::Command};

const SOURCES: &[&str] = &[
    "qz_ctx.cu", "qz_api.cu", "row_scan.cu", "row_scan_dot.cu", "row_scan_l1.cu", "row_scan_xor.cu", "point_kernels.cu",
    "pq.cu", "pq_scan2.cu", "encode.cu", "encode_api.cu", "kmeans.cu", "pq_encode_api.cu", "u8_batch.cu", "peer.cu",
    "chunked.cu", "no_alloc.cu",
];
// the flags of quantization_b200/build.py: sm_100a only, no FMA contraction (the scorers restate the reference's
// separately rounded multiply and add, encoded_vectors_u8.rs:347)
const FLAGS: &[&str] = &[
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler",
    "-fPIC,-fvisibility=hidden,-O2", "--fmad=false",
];

fn main() {
    if env::var("CARGO_FEATURE_CUDA").is_err() {
        build_simd_utils(); // the existing cc::Build over cpp/*.c, unchanged (quantization/build.rs:18-40)
        return;
    }
    println!("cargo:rerun-if-env-changed=QZ_CUDA_DIR");
    println!("cargo:rerun-if-env-changed=QZ_CUDA_SRC");
    println!("cargo:rerun-if-env-changed=NVCC");
    if let Ok(dir) = env::var("QZ_CUDA_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
    } else {
        let src = PathBuf::from(env::var("QZ_CUDA_SRC").expect("set QZ_CUDA_DIR (prebuilt library) or QZ_CUDA_SRC (csrc/)"));
        let out = PathBuf::from(env::var("OUT_DIR").unwrap());
        let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
        let mut objects = Vec::new();
        for f in SOURCES {
            let cu = src.join(f);
            println!("cargo:rerun-if-changed={}", cu.display());
            let obj = out.join(f.replace(".cu", ".o"));
            let ok = Command::new(&nvcc).args(FLAGS).arg("-c").arg(&cu).arg("-o").arg(&obj).status().expect("nvcc not found").success();
            assert!(ok, "nvcc failed on {f}");
            objects.push(obj);
        }
        let lib = out.join("libqz_cuda.so");
        let ok = Command::new(&nvcc).arg("-shared").arg("-o").arg(&lib).args(&objects)
            .args(["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "shared"]).status().unwrap().success();
        assert!(ok, "linking libqz_cuda.so failed");
        println!("cargo:rustc-link-search=native={}", out.display());
    }
    println!("cargo:rustc-link-lib=dylib=qz_cuda");
    println!("cargo:rustc-link-lib=dylib=cudart");
}

fn build_simd_utils() {
    // quantization/build.rs:18-40 as it stands today
}
