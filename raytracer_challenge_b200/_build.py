"""In-tree build of the native libraries (nvcc for sm_100a; g++ for the harness glue and the oracle).

Outputs (git-ignored, shipped to the GPU box by gpurun):
  raytracer_challenge_b200/lib/librtc.so         product: CUDA kernels + C ABI of include/rtc.h
  raytracer_challenge_b200/lib/librtc_driver.so  harness glue: named scenes -> host builder API -> C ABI
  build/liboracle.so, build/test_oracle          test infrastructure (oracle/), built via oracle/Makefile
"""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "raytracer_challenge_b200", "csrc")
LIB = os.path.join(ROOT, "raytracer_challenge_b200", "lib")
OBJ = os.path.join(ROOT, "build", "obj")

NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# -fmad=false: no multiply-add contraction, so f64 results are bit-identical to the reference's scalar code
DEVICE_FLAGS = ["-std=c++17", "-O3", "-lineinfo", "-fmad=false", "-Xcompiler", "-fPIC", "-Wno-deprecated-gpu-targets"] + ARCH + os.environ.get("RTC_NVCC_EXTRA", "").split()
HOST_FLAGS = ["-std=c++17", "-O2", "-Xcompiler", "-fPIC,-Wall", "-Wno-deprecated-gpu-targets"] + ARCH


def read_units():
    """(file, kind) pairs of csrc/UNITS — the one list of translation units, shared with rust/build.rs."""
    units = []
    with open(os.path.join(CSRC, "UNITS")) as f:
        for line in f:
            line = line.split("#")[0].split()
            if len(line) == 2:
                units.append((line[0], line[1]))
    return units


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, **kw):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, **kw)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + "\n")
        raise RuntimeError("build failed: " + " ".join(cmd[:3]))
    return r.stdout


def build_rtc(force=False, verbose=False):
    os.makedirs(LIB, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".hpp", ".cuh"))]
    headers.append(os.path.join(ROOT, "include", "rtc.h"))
    cuda_flags = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC", "-Wno-deprecated-gpu-targets"] + ARCH
    flags_of = {"device": DEVICE_FLAGS + (["-Xptxas", "-v"] if verbose else []), "cuda": cuda_flags, "host": HOST_FLAGS}
    units = [(name, flags_of[kind]) for name, kind in read_units()]
    objs = []
    log = ""
    for name, flags in units:
        src = os.path.join(CSRC, name)
        obj = os.path.join(OBJ, name + ".o")
        objs.append(obj)
        if force or _newer(obj, [src] + headers):
            log += _run([NVCC] + flags + ["-x", "cu", "-c", src, "-o", obj])
    so = os.path.join(LIB, "librtc.so")
    if force or _newer(so, objs):
        _run([NVCC, "-shared", "-o", so] + objs + ["-Wno-deprecated-gpu-targets"] + ARCH + ["-cudart", "shared"])
    return so, log


def build_driver(force=False):
    so = os.path.join(LIB, "librtc_driver.so")
    srcs = [os.path.join(ROOT, "driver", "driver.cpp"), os.path.join(ROOT, "raytracer_challenge_b200", "host", "raytracer.hpp"),
            os.path.join(ROOT, "scenes", "scenes.hpp"), os.path.join(ROOT, "include", "rtc.h")]
    if force or _newer(so, srcs + [os.path.join(LIB, "librtc.so")]):
        _run(["g++", "-std=c++17", "-O2", "-fPIC", "-Wall", "-shared", "-o", so, srcs[0], "-L" + LIB, "-lrtc", "-Wl,-rpath,$ORIGIN"])
    return so


def build_oracle(force=False):
    out = os.path.join(ROOT, "build")
    cmd = ["make", "-C", os.path.join(ROOT, "oracle"), "OUT=" + out]
    if force:
        cmd.append("-B")
    _run(cmd)
    return os.path.join(out, "liboracle.so")


def build_all(force=False, verbose=False):
    so, log = build_rtc(force, verbose)
    drv = build_driver(force)
    orc = build_oracle(force)
    return {"librtc": so, "driver": drv, "oracle": orc, "log": log}


if __name__ == "__main__":
    res = build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    if res["log"]:
        print(res["log"])
    print({k: v for k, v in res.items() if k != "log"})
