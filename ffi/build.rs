// Links librtcore.so.  RTCORE_LIB_DIR = directory holding it (default: ../rustraytracer_b200, where `make -C rustraytracer_b200/csrc` leaves it).
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("RTCORE_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("..").join("rustraytracer_b200")
    });
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=rtcore");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=RTCORE_LIB_DIR");
    println!("cargo:rerun-if-changed=../include/rtcore.h");
}
