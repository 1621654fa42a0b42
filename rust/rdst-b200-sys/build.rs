// Links libb200sort.so (built by `make -C rdst_b200/csrc` in the engine's repository).
// B200SORT_LIB_DIR = directory holding the library; default: ../../rdst_b200 relative to this crate.
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("B200SORT_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../rdst_b200")
    });
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=b200sort");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=B200SORT_LIB_DIR");
    println!("cargo:rerun-if-changed=../../include/b200sort.h");
}
