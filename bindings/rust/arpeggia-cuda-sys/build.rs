// Link against the in-tree library: ARPEGGIA_B200_LIB_DIR=<repo>/arpeggia_b200/_build cargo build
fn main() {
    if let Ok(dir) = std::env::var("ARPEGGIA_B200_LIB_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
    }
    println!("cargo:rustc-link-lib=dylib=arpeggia_b200");
    println!("cargo:rerun-if-env-changed=ARPEGGIA_B200_LIB_DIR");
}
