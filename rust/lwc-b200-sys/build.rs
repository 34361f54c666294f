// Links liblwc_b200.so.  LWC_B200_LIB_DIR points at the directory holding the library
// (llm-weighted-consensus_b200/ in this repository after `python __graft_entry__.py`).
fn main() {
    if let Ok(dir) = std::env::var("LWC_B200_LIB_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
    }
    println!("cargo:rustc-link-lib=dylib=lwc_b200");
    println!("cargo:rerun-if-env-changed=LWC_B200_LIB_DIR");
}
