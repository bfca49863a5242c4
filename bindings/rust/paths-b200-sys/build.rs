// Links libpaths_b200.so.  PATHS_B200_LIB_DIR overrides the in-tree location (paths_b200/lib after `make -C paths_b200`).
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("PATHS_B200_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../../paths_b200/lib")
    });
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=paths_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=PATHS_B200_LIB_DIR");
}
