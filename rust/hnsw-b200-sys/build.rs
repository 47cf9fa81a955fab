// UNCOMPILED (no Rust toolchain in the build image) -- see rust/README.md.
//
// Either link a prebuilt libhnswb200.so (HNSWB200_LIB_DIR), or compile the CUDA sources with the `cc` crate for
// sm_100a, as BASELINE.json's north star describes ("thin C-ABI FFI crate: cc-built .cu plus extern \"C\"").
use std::{env, path::PathBuf};

fn main() {
    println!("cargo:rerun-if-env-changed=HNSWB200_LIB_DIR");
    if let Ok(dir) = env::var("HNSWB200_LIB_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-lib=dylib=hnswb200");
        println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
        return;
    }
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    let csrc = root.join("hnsw_itu_b200/csrc");
    let sources = [
        "api.cu", "index.cu", "search.cu", "bruteforce.cu", "gt_tcgen05.cu", "gt_mxf4.cu", "gt_peak.cu", "synth.cu",
        "builder.cu", "multi.cu",
    ];
    let mut b = cc::Build::new();
    b.cuda(true)
        .cudart("shared")
        .flag("-gencode")
        .flag("arch=compute_100a,code=sm_100a")
        .flag("-lineinfo")
        .flag("-O3")
        .flag("-std=c++17")
        .include(root.join("include"));
    for s in sources {
        b.file(csrc.join(s));
        println!("cargo:rerun-if-changed={}", csrc.join(s).display());
    }
    // idx_format.cpp is plain C++ but includes cuda_runtime.h through hb_internal.h: compile it as CUDA too
    b.flag("-x").flag("cu").file(csrc.join("idx_format.cpp"));
    b.compile("hnswb200");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=dl"); // NCCL is resolved with dlopen at first multi-GPU use
}
