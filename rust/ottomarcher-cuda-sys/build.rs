// Builds libottomarcher_cuda for sm_100a with nvcc and links it.  Mirrors
// raytracingoneweekend_b200/build.py (per-file flags matter: the parity kernels are -fmad=false).
// NOT COMPILED IN THIS IMAGE (no rustc/cargo).
use std::{env, path::PathBuf, process::Command};

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    let csrc = root.join("raytracingoneweekend_b200/csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "nvcc".into());
    let common = ["-O3", "-std=c++17", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
                  "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-fvisibility=hidden"];
    // the same list and per-file flags as raytracingoneweekend_b200/build.py SOURCES (tests/test_rust_ffi.py compares the two)
    let sources: [(&str, &[&str]); 7] = [
        ("om_kernels_fast.cu", &[]), ("om_kernels_exact.cu", &["-fmad=false"]), ("om_api.cu", &["-fmad=false"]), ("om_lbvh.cu", &[]),
        ("om_host.cpp", &[]), ("om_scene.cpp", &[]), ("om_bvh.cpp", &[]),
    ];
    let mut objs = Vec::new();
    for (src, extra) in sources.iter() {
        let obj = out.join(format!("{}.o", src));
        let ok = Command::new(&nvcc).args(&common).args(*extra).arg("-c").arg(csrc.join(src)).arg("-o").arg(&obj)
            .status().expect("nvcc not found").success();
        assert!(ok, "nvcc failed for {}", src);
        objs.push(obj);
        println!("cargo:rerun-if-changed={}", csrc.join(src).display());
    }
    let lib = out.join("libottomarcher_cuda.so");
    let ok = Command::new(&nvcc).args(&["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o"]).arg(&lib).args(&objs)
        .status().unwrap().success();
    assert!(ok, "link failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=ottomarcher_cuda");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rerun-if-changed={}", root.join("include/ottomarcher.h").display());
}
