// build.rs — compiles the B200 backend (ray-tracer_b200/csrc + host/host_capi.cpp) into a static library and links it
// into the `ray-tracer` crate.  Drop this file at the root of aiifabbf/ray-tracer (next to Cargo.toml), copy or
// sub-module this repository to `./b200/`, and add to Cargo.toml:
//
//     [package]
//     build = "build.rs"
//     links = "rt_b200"
//
// Inputs are EXACTLY the objects of ray-tracer_b200/Makefile (`OBJ := build/kernels.o build/capi.o build/lbvh.o
// build/bvh_build.o build/host_capi.o`); tests/test_rust_binding.py diffs this list against the Makefile.  nvcc is
// asked for sm_100a only (no PTX for other targets, no fallback): `-gencode arch=compute_100a,code=sm_100a`.
// Link line: cudart (static, as `nvcc -cudart static` does in the Makefile), z (PNG / inflate in host/image_io.hpp),
// dl (NCCL is resolved with dlopen by rt_init_multi — the backend has NO link-time dependency on libnccl, see
// include/rt_b200.h), stdc++, pthread, rt.
//
// Environment: CUDA_HOME (default /usr/local/cuda), RT_B200_DIR (default ./b200), NVCC / CXX override the compilers.
// No Rust toolchain exists in the image this backend was developed in, so this file has never been run by its
// authors; every path and flag below is the Makefile's.

use std::env;
use std::path::{Path, PathBuf};
use std::process::Command;

/// (source relative to ray-tracer_b200/, object name, compiled by nvcc?)
const UNITS: &[(&str, &str, bool)] = &[
    ("csrc/kernels.cu", "kernels.o", true),
    ("csrc/capi.cu", "capi.o", true),
    ("csrc/lbvh.cu", "lbvh.o", true),
    ("csrc/bvh_build.cpp", "bvh_build.o", false),
    ("host/host_capi.cpp", "host_capi.o", false),
];

/// Headers whose change must trigger a rebuild (the Makefile's HDR list).
const HEADERS: &[&str] = &[
    "csrc/box_prims.h",
    "csrc/bulk_copy.cuh",
    "csrc/device_scene.h",
    "csrc/intersect.cuh",
    "csrc/shade.cuh",
    "csrc/kernels.h",
    "csrc/philox.h",
    "csrc/bvh_build.h",
    "csrc/lbvh.h",
    "host/math.hpp",
    "host/scene_graph.hpp",
    "host/flatten.hpp",
    "host/gpu.hpp",
    "host/image_io.hpp",
    "host/jpeg_decode.hpp",
    "host/api_bundle.hpp",
    "examples/scenes.hpp",
    "../include/rt_b200.h",
    "../include/rt_b200_host.h",
];

fn run(cmd: &mut Command) {
    let status = cmd.status().unwrap_or_else(|e| panic!("cannot run {:?}: {}", cmd, e));
    if !status.success() {
        panic!("{:?} failed with {}", cmd, status);
    }
}

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let root = PathBuf::from(env::var("RT_B200_DIR").unwrap_or_else(|_| "b200".to_string()));
    let pkg = root.join("ray-tracer_b200");
    let cuda = PathBuf::from(env::var("CUDA_HOME").unwrap_or_else(|_| "/usr/local/cuda".to_string()));
    let nvcc = env::var("NVCC").unwrap_or_else(|_| cuda.join("bin/nvcc").to_string_lossy().into_owned());
    let cxx = env::var("CXX").unwrap_or_else(|_| "g++".to_string());

    let mut objects: Vec<PathBuf> = Vec::new();
    for (src, obj, is_cuda) in UNITS {
        let src_path = pkg.join(src);
        let obj_path = out.join(obj);
        println!("cargo:rerun-if-changed={}", src_path.display());
        if *is_cuda {
            // NVFLAGS of the Makefile (without -Xptxas -v: that is for the register / spill log the authors read)
            run(Command::new(&nvcc)
                .args(&["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17"])
                .args(&["-Xcompiler", "-fPIC,-Wall,-Wno-unknown-pragmas"])
                .arg("-c")
                .arg(&src_path)
                .arg("-o")
                .arg(&obj_path));
        } else {
            // CXXFLAGS of the Makefile; -ffp-contract=off: bound() of the host mirror must round exactly like the reference's f64 code
            run(Command::new(&cxx)
                .args(&["-O2", "-std=c++17", "-fPIC", "-Wall", "-Wextra", "-Wno-unused-parameter", "-Wno-unknown-pragmas", "-ffp-contract=off"])
                .arg(format!("-I{}", cuda.join("include").display()))
                .arg("-c")
                .arg(&src_path)
                .arg("-o")
                .arg(&obj_path));
        }
        objects.push(obj_path);
    }
    for h in HEADERS {
        println!("cargo:rerun-if-changed={}", pkg.join(h).display());
    }

    // device-link step: none needed (no relocatable device code: every kernel and its callers live in one translation unit each)
    let lib = out.join("librt_b200.a");
    let _ = std::fs::remove_file(&lib);
    run(Command::new("ar").arg("crs").arg(&lib).args(&objects));

    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=static=rt_b200");
    let cuda_lib: &Path = &cuda.join("lib64");
    println!("cargo:rustc-link-search=native={}", cuda_lib.display());
    println!("cargo:rustc-link-lib=static=cudart_static");
    for l in &["z", "dl", "rt", "pthread", "stdc++"] {
        println!("cargo:rustc-link-lib=dylib={}", l);
    }
    println!("cargo:rerun-if-env-changed=CUDA_HOME");
    println!("cargo:rerun-if-env-changed=RT_B200_DIR");
}
