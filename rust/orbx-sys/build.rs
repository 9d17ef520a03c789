// Compiles the hand-written CUDA kernels of liborbx for sm_100a and links them.
// NOT compiled in the build image of this repository (no cargo/rustc there) — shipped as source.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    let csrc = root.join("slam-rs_b200/csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    let mut objs = vec![];
    // the source list is the Makefile's (one place to add a file): "SRCS := api.cu match.cu ..."
    let mk = std::fs::read_to_string(csrc.join("Makefile")).expect("csrc/Makefile");
    let srcs: Vec<String> = mk.lines().find(|l| l.starts_with("SRCS")).expect("SRCS line")
        .split(":=").nth(1).unwrap().split_whitespace().map(|f| f.trim_end_matches(".cu").to_string()).collect();
    println!("cargo:rerun-if-changed={}", csrc.join("Makefile").display());
    for f in srcs.iter() {
        let src = csrc.join(format!("{f}.cu"));
        let obj = out.join(format!("{f}.o"));
        let st = Command::new(&nvcc)
            .args(["-O3", "-std=c++17", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
                   "-Xcompiler", "-fPIC", "-c"])
            .arg(&src).arg("-o").arg(&obj).status().expect("nvcc not found");
        assert!(st.success(), "nvcc failed on {}", src.display());
        println!("cargo:rerun-if-changed={}", src.display());
        objs.push(obj);
    }
    let lib = out.join("liborbx.a");
    let st = Command::new("ar").arg("rcs").arg(&lib).args(&objs).status().unwrap();
    assert!(st.success());
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=static=orbx");
    println!("cargo:rustc-link-search=native=/usr/local/cuda/lib64");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=dl");  // csrc/comm.cu binds NCCL at run time (dlopen of libnccl.so.2)
    println!("cargo:rustc-link-lib=dylib=stdc++");
}
