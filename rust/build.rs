// Links liblinalg_b200.so.  LINALG_B200_DIR points at the directory that holds it (default: the in-tree
// build output, ../linear_algebra_b200, made by `make -C linear_algebra_b200/csrc`).
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("LINALG_B200_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("..").join("linear_algebra_b200")
    });
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=linalg_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=LINALG_B200_DIR");
}
