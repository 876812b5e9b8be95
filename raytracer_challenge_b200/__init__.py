"""B200-native render path for pacellie/raytracer_challenge (per-pixel render loop only).

The product is native: CUDA kernels + a C ABI (include/rtc.h, built into lib/librtc.so) and a C++ host
builder API mirroring the reference crate (host/raytracer.hpp).  This Python package is harness plumbing:
it builds the libraries in-tree and exposes thin ctypes bindings for tests/, bench.py and __graft_entry__.py.
"""
from .binding import (  # noqa: F401
    RtcError,
    Opts,
    Stats,
    Session,
    lib_paths,
    load_rtc,
    load_driver,
    quantize,
    shard_pixels,
    device_count,
    pinned_array,
    free_pinned,
    microbench,
    par_render_multi,
    set_tuning,
    tuning,
    Tuning,
    BVH_DEVICE,
    BVH_HOST,
    FLAG_BRUTE_FORCE,
    FLAG_PROFILE,
    FLAG_SCATTER_ROWS,
    SCENES,
    DATA_DIR,
)
