// Builds (if needed) and links libm4ri_b200.so. Replaces the autoreconf/configure/make of the
// vendored M4RI that m4ri-sys/build.rs:27-69 runs, for every symbol m4ri-rust references
// (one allocator: never link a second libm4ri next to it).
use std::{env, path::PathBuf, process::Command};

fn main() {
    // checkout of the m4ri-b200 repository; defaults to two levels above this crate
    let root: PathBuf = env::var("M4RI_B200_DIR")
        .map(PathBuf::from)
        .unwrap_or_else(|_| PathBuf::from(env!("CARGO_MANIFEST_DIR")).join("../.."));
    let lib = root.join("m4ri-rust_b200/lib");
    if !lib.join("libm4ri_b200.so").exists() {
        let status = Command::new("python")
            .arg(root.join("m4ri-rust_b200/build.py"))
            .status()
            .expect("cannot run m4ri-rust_b200/build.py (needs nvcc with sm_100a support)");
        assert!(status.success(), "building libm4ri_b200.so failed");
    }
    println!("cargo:rustc-link-search=native={}", lib.display());
    println!("cargo:rustc-link-lib=dylib=m4ri_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", lib.display());
    println!("cargo:rerun-if-env-changed=M4RI_B200_DIR");
}
