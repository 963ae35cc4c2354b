// build.rs — compile the CUDA core for sm_100a and link it statically (UNCOMPILED HERE: no Rust toolchain in the image).
// Expects the kernel sources of this repo under spt/csrc and the header under spt/include.
use std::process::Command;

fn main() {
    let out = std::env::var("OUT_DIR").unwrap();
    let src = ["spt_api.cu", "trav_kernels.cu", "shade_a.cu", "shade_b.cu", "shade_c.cu", "shade_a_lean.cu", "shade_b_lean.cu", "shade_c_lean.cu", "misc_kernels.cu", "bvh_build_gpu.cu", "flatten.cpp", "bvh_build.cpp"];
    let mut objs = vec![];
    for s in src {
        let o = format!("{}/{}.o", out, s);
        let st = Command::new("nvcc")
            .args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-prec-div=false", "-prec-sqrt=false",
                   "-Xcompiler", "-fPIC,-ffp-contract=off", "-c", &format!("spt/csrc/{}", s), "-o", &o])
            .status()
            .expect("nvcc not found");
        assert!(st.success(), "nvcc failed on {}", s);
        println!("cargo:rerun-if-changed=spt/csrc/{}", s);
        objs.push(o);
    }
    // device code is fully resolved per translation unit (no relocatable device code), so a plain archive is enough
    let st = Command::new("ar").arg("crs").arg(format!("{}/libspt_cuda.a", out)).args(&objs).status().unwrap();
    assert!(st.success());
    println!("cargo:rustc-link-search=native={}", out);
    println!("cargo:rustc-link-lib=static=spt_cuda");
    println!("cargo:rustc-link-search=native=/usr/local/cuda/lib64");
    println!("cargo:rustc-link-lib=cudart");
    println!("cargo:rustc-link-lib=stdc++");
    println!("cargo:rustc-link-lib=dl"); // NCCL is dlopen()ed by the library (spt_comm_*), not linked
}
