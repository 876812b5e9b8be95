// build.rs for the `raytracer` crate (reference: /root/reference/Cargo.toml has no build script today).
// UNBUILT IN THIS REPOSITORY'S IMAGE: no cargo / rustc there.  Shipped as the binding a maintainer would add.
//
// Builds (or locates) librtc.so — the CUDA kernels + C ABI of include/rtc.h — and links it.
//   RTC_LIB_DIR=<dir>   use a prebuilt librtc.so (e.g. raytracer_challenge_b200/lib after `python -m raytracer_challenge_b200._build`)
//   otherwise           compile csrc/ with nvcc for sm_100a into OUT_DIR
use std::env;
use std::path::PathBuf;
use std::process::Command;

fn main() {
    println!("cargo:rerun-if-env-changed=RTC_LIB_DIR");
    if let Ok(dir) = env::var("RTC_LIB_DIR") {
        println!("cargo:rustc-link-search=native={}", dir);
        println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir);
        println!("cargo:rustc-link-lib=dylib=rtc");
        return;
    }
    let root = PathBuf::from(env::var("RTC_SOURCE_DIR").expect("set RTC_SOURCE_DIR to the raytracer_challenge_b200 checkout or RTC_LIB_DIR to a prebuilt librtc.so"));
    let csrc = root.join("raytracer_challenge_b200").join("csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "nvcc".into());
    let arch = ["-gencode", "arch=compute_100a,code=sm_100a"];
    // The translation units and their flag classes come from csrc/UNITS, the list the Python build reads too.
    //   device: -fmad=false (no multiply-add contraction: f64 results stay bit-identical to this crate's scalar code)
    let listing = std::fs::read_to_string(csrc.join("UNITS")).expect("cannot read csrc/UNITS");
    println!("cargo:rerun-if-changed={}", csrc.join("UNITS").display());
    let mut units: Vec<(String, Vec<&str>)> = Vec::new();
    for line in listing.lines() {
        let line = line.split('#').next().unwrap();
        let mut it = line.split_whitespace();
        if let (Some(name), Some(kind)) = (it.next(), it.next()) {
            let flags: Vec<&str> = match kind {
                "device" => vec!["-O3", "-lineinfo", "-fmad=false"],
                "cuda" => vec!["-O3", "-lineinfo"],
                "host" => vec!["-O2"],
                other => panic!("csrc/UNITS: unknown kind {}", other),
            };
            units.push((name.to_string(), flags));
        }
    }
    let mut objs = Vec::new();
    for (name, flags) in units.iter() {
        let obj = out.join(format!("{}.o", name));
        let status = Command::new(&nvcc)
            .args(["-std=c++17", "-Xcompiler", "-fPIC", "-x", "cu", "-c"])
            .args(arch)
            .args(flags.iter())
            .arg(csrc.join(name.as_str()))
            .arg("-o")
            .arg(&obj)
            .status()
            .expect("failed to run nvcc");
        assert!(status.success(), "nvcc failed on {}", name);
        println!("cargo:rerun-if-changed={}", csrc.join(name.as_str()).display());
        objs.push(obj);
    }
    let so = out.join("librtc.so");
    let status = Command::new(&nvcc).arg("-shared").arg("-o").arg(&so).args(&objs).args(arch).args(["-cudart", "shared"]).status().expect("failed to run nvcc");
    assert!(status.success(), "linking librtc.so failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", out.display());
    println!("cargo:rustc-link-lib=dylib=rtc");
}
