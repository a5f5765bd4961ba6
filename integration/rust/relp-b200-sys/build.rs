// Builds librelp_b200.so with nvcc for sm_100a (the only target) and links it.  RELP_B200_ROOT = checkout of this
// repository (default: three directories up).
fn main() {
    let root = std::env::var("RELP_B200_ROOT").unwrap_or_else(|_| "../../..".into());
    let out = std::path::PathBuf::from(std::env::var("OUT_DIR").unwrap());
    let so = out.join("librelp_b200.so");
    let src = format!("{}/rust-lp_b200/csrc/relp_capi.cu", root);
    let src2 = format!("{}/rust-lp_b200/csrc/relp_fastloop.cu", root);      // the loop kernel without the arbitrary-length tier
    println!("cargo:rerun-if-changed={}", src);
    println!("cargo:rerun-if-changed={}", src2);
    println!("cargo:rerun-if-changed={}/rust-lp_b200/csrc/relp_wcb.cuh", root);
    println!("cargo:rerun-if-changed={}/rust-lp_b200/csrc/relp_kernels.cuh", root);
    println!("cargo:rerun-if-changed={}/rust-lp_b200/csrc/relp_arith.cuh", root);
    println!("cargo:rerun-if-changed={}/include/relp_b200.h", root);
    let st = std::process::Command::new(std::env::var("NVCC").unwrap_or_else(|_| "nvcc".into()))
        .args(&["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
                "-Xcompiler", "-fPIC", "-shared"])
        .arg(&src).arg(&src2).arg("-o").arg(&so).arg("-ldl")
        .status().expect("nvcc not found");
    assert!(st.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=relp_b200");
}
