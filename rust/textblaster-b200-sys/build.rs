// NOT COMPILED IN THIS REPOSITORY (no Rust toolchain in the build image).
// Points the linker at libtextblaster_b200.so: set TEXTBLASTER_B200_LIB_DIR to the directory that holds it
// (the repository's textblaster_b200/ after `python -c 'import __graft_entry__ as g; g.build()'`).
fn main() {
    if let Ok(dir) = std::env::var("TEXTBLASTER_B200_LIB_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
    }
    println!("cargo:rustc-link-lib=dylib=textblaster_b200");
    println!("cargo:rerun-if-env-changed=TEXTBLASTER_B200_LIB_DIR");
}
