// Links libb200img.so (built by `python -m image_b200.build` at the repository root).
fn main() {
    let dir = std::env::var("B200IMG_LIB_DIR").unwrap_or_else(|_| "../../image_b200/lib".to_string());
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=b200img");
    println!("cargo:rerun-if-env-changed=B200IMG_LIB_DIR");
}
