// Links the prebuilt libctg_groth16.so (make -C contangle-zkcp_b200/csrc).  CTG_LIB_DIR overrides the search path.
fn main() {
    let dir = std::env::var("CTG_LIB_DIR").unwrap_or_else(|_| format!("{}/../..", env!("CARGO_MANIFEST_DIR")));
    println!("cargo:rustc-link-search=native={}", dir);
    println!("cargo:rustc-link-lib=dylib=ctg_groth16");
    println!("cargo:rerun-if-env-changed=CTG_LIB_DIR");
}
