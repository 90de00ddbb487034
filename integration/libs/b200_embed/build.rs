// Points rustc at libb200_embed.so (built by codename-symbiont_b200/build.py with nvcc for sm_100a).
fn main() {
    let dir = std::env::var("B200_EMBED_LIB_DIR").unwrap_or_else(|_| "/usr/local/lib".into());
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=b200_embed");
    println!("cargo:rerun-if-env-changed=B200_EMBED_LIB_DIR");
}
