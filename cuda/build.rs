// build.rs of the `gffx-cuda` crate: drives nvcc exactly as this repository's Makefile does (same source lists — kept in
// step by tests/test_abi.py::test_rust_crate_sources_match_the_makefile) and links the static library into the `gffx`
// binary.  `native/` = this repository's gffx_b200/csrc + include/ (git subtree or symlink).  Not compiled in the build
// image (no cargo / rustc there); see INTEGRATION.md.
use std::{env, path::PathBuf, process::Command};
fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    let src = "native/csrc";
    let cu = ["engine.cu", "counts.cu", "bedparse.cu", "linescan.cu"];
    let cpp = ["util.cpp", "winjoin_build.cpp", "counts_index.cpp", "host_index.cpp", "linecache.cpp", "reads.cpp", "index_builder.cpp", "synth.cpp", "capi.cpp"];
    let mut objs = vec![];
    for f in cu.iter().chain(cpp.iter()) {
        let o = out.join(format!("{f}.o"));
        let st = Command::new(&nvcc)
            .args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
                   "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Inative/include", "-c"])
            .arg(format!("{src}/{f}")).arg("-o").arg(&o).status().expect("nvcc");
        assert!(st.success(), "nvcc failed on {f}");
        objs.push(o);
    }
    let lib = out.join("libgffx_b200.a");
    assert!(Command::new("ar").arg("crs").arg(&lib).args(&objs).status().unwrap().success());
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=static=gffx_b200");
    println!("cargo:rustc-link-search=native=/usr/local/cuda/lib64");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=stdc++");
    println!("cargo:rerun-if-changed=native");
}
