// Compiles the hand-written CUDA of librfgpu for sm_100a and links it (north_star: "a new CUDA crate whose build.rs compiles
// hand-written .cu for sm_100a and exposes a thin extern \"C\" FFI").  RFGPU_CSRC points at rdf_fusion_b200/csrc.
fn main() {
    let csrc = std::env::var("RFGPU_CSRC").unwrap_or_else(|_| "csrc".to_string());
    let sources = [
        "ctx.cu", "table.cu", "filter.cu", "scan.cu", "join.cu", "shuffle.cu", "aggregate.cu", "sort.cu", "dict.cu", "synth.cu",
        "scan_plan.cpp",
    ];
    let mut b = cc::Build::new();
    b.cuda(true)
        .flag("-gencode").flag("arch=compute_100a,code=sm_100a")
        .flag("-O3").flag("-std=c++17").flag("-lineinfo").flag("--expt-relaxed-constexpr")
        .include(format!("{csrc}/../../include"));
    for s in sources {
        b.file(format!("{csrc}/{s}"));
        println!("cargo:rerun-if-changed={csrc}/{s}");
    }
    b.compile("rfgpu");
    println!("cargo:rustc-link-lib=cudart");
    println!("cargo:rustc-link-lib=dl"); // NCCL is dlopen()ed at run time (shuffle.cu)
}
