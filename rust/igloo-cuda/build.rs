// Builds csrc/ (a copy or submodule of igloo_b200/csrc + include/) with nvcc for sm_100a and
// links the resulting shared library.  NVCC can be overridden through the environment.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let lib = out.join("libigloo_cuda.so");
    let status = Command::new("make")
        .arg("-C").arg(manifest.join("csrc"))
        .arg("-j8")
        .arg(format!("BUILD={}", out.join("obj").display()))
        .arg(format!("OUT={}", lib.display()))
        .status()
        .expect("make not found");
    assert!(status.success(), "nvcc build of libigloo_cuda.so (-gencode arch=compute_100a,code=sm_100a) failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=igloo_cuda");
    println!("cargo:rerun-if-changed=csrc");
    println!("cargo:rerun-if-changed=include");
}
