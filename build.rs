// build.rs of the vgen crate once src/gpu.rs binds libvgen_cuda.so (include/vgen_cuda.h) instead of wgpu.
//
// The CUDA library is built by `make -C vgen_b200/csrc` (nvcc -gencode arch=compute_100a,code=sm_100a); this script only
// tells cargo where to find it. VGEN_CUDA_LIB_DIR overrides the in-tree location; VGEN_CUDA_BUILD=1 runs the Makefile first.
use std::env;
use std::path::PathBuf;
use std::process::Command;

fn main() {
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").expect("CARGO_MANIFEST_DIR"));
    let lib_dir = env::var("VGEN_CUDA_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| manifest.join("vgen_b200"));
    if env::var("VGEN_CUDA_BUILD").as_deref() == Ok("1") {
        let status = Command::new("make")
            .arg("-C")
            .arg(manifest.join("vgen_b200").join("csrc"))
            .arg("-j")
            .status()
            .expect("failed to run make for libvgen_cuda.so");
        assert!(status.success(), "building libvgen_cuda.so failed");
    }
    assert!(
        lib_dir.join("libvgen_cuda.so").exists(),
        "libvgen_cuda.so not found in {} (run `make -C vgen_b200/csrc` or set VGEN_CUDA_LIB_DIR); there is no CPU fallback",
        lib_dir.display()
    );
    println!("cargo:rustc-link-search=native={}", lib_dir.display());
    println!("cargo:rustc-link-lib=dylib=vgen_cuda");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", lib_dir.display());
    println!("cargo:rerun-if-env-changed=VGEN_CUDA_LIB_DIR");
    println!("cargo:rerun-if-changed=include/vgen_cuda.h");
}
