// Links the prebuilt shared library (make -C dsd2dxd_b200/csrc).  DSD2PCM_B200_LIB_DIR = directory holding
// libdsd2pcm_b200.so; defaults to the in-tree location relative to this crate.
fn main() {
    let dir = std::env::var("DSD2PCM_B200_LIB_DIR").unwrap_or_else(|_| {
        let here = std::env::var("CARGO_MANIFEST_DIR").unwrap();
        format!("{here}/../../dsd2dxd_b200")
    });
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=dsd2pcm_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
    println!("cargo:rerun-if-env-changed=DSD2PCM_B200_LIB_DIR");
}
