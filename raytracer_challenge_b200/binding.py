"""ctypes bindings over include/rtc.h (librtc.so) and the scene harness (librtc_driver.so).

No CPU fallback: if the CUDA library is missing or no device is visible, rendering raises RtcError.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_DIR = os.environ.get("RTC_LIB_DIR") or os.path.join(ROOT, "raytracer_challenge_b200", "lib")  # (the variable: compile-time variants side by side, tools/build_variant.sh)
DATA_DIR = os.path.join(ROOT, "data")
SCENES = ["default_world", "three_spheres", "glass_air_bubble", "hexagon", "benchmark", "teapot", "teapot_low", "csg_noise",
          "chapter11_title", "chapter12_title", "chapter13_title", "chapter14_title", "cover", "cone_slope", "axis_aligned", "no_shadow_flat", "many_gates", "glass_shells", "tri_ghost", "csg_cuts"]


class RtcError(RuntimeError):
    pass


class Opts(C.Structure):  # rtc_opts
    _fields_ = [("fuel", C.c_int32), ("spp_x", C.c_int32), ("spp_y", C.c_int32), ("shard_index", C.c_int32),
                ("shard_count", C.c_int32), ("tile_w", C.c_int32), ("tile_h", C.c_int32), ("flags", C.c_uint32),
                ("win_x0", C.c_int32), ("win_y0", C.c_int32), ("win_w", C.c_int32), ("win_h", C.c_int32)]

    def __init__(self, fuel=5, spp=(1, 1), shard=(0, 1), tile=(0, 0), flags=0, window=None):
        super().__init__()
        self.fuel = fuel
        self.spp_x, self.spp_y = spp
        self.shard_index, self.shard_count = shard
        self.tile_w, self.tile_h = tile
        self.flags = flags
        if window:
            self.win_x0, self.win_y0, self.win_w, self.win_h = window


class Stats(C.Structure):  # rtc_stats
    _fields_ = [("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_reflected", C.c_uint64),
                ("rays_refracted", C.c_uint64), ("rays_ref", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("pixels", C.c_uint64), ("ms_total", C.c_double), ("ms_kernels", C.c_double),
                ("ms_trace_shade", C.c_double), ("ms_d2h", C.c_double), ("overflow", C.c_int32), ("reserved", C.c_int32),
                ("surface_hits", C.c_uint64), ("bvh_node_visits", C.c_uint64), ("triangle_tests", C.c_uint64),
                ("primitive_tests", C.c_uint64), ("ray_transforms", C.c_uint64),
                ("ms_trace", C.c_double), ("ms_shadow", C.c_double), ("ms_shade", C.c_double), ("max_node_visits", C.c_uint64),
                ("shadow_queries_culled", C.c_uint64)]

    @property
    def rays(self):
        return self.rays_primary + self.rays_shadow + self.rays_reflected + self.rays_refracted

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


class Tuning(C.Structure):  # rtc_tuning
    _fields_ = [(n, C.c_int32) for n in ("batches", "batch_samples", "pipes", "blocks_per_sm", "no_taper", "no_copy_stream", "no_stream_priorities",
                                           "no_graphs", "no_fast_planes", "no_plain_shade", "two_pass", "lbvh_leaf", "bvh_builder", "build_threads",
                                           "sah_leaf_weight_x100", "sah_bins", "sah_small_node", "host_paced_copies")] + [("reserved", C.c_int32 * 14)]


BVH_DEVICE, BVH_HOST = 1, 2  # rtc_tuning.bvh_builder
# Harness convenience (tools/, sweeps): these environment variables are translated into rtc_set_tuning when the library is
# loaded.  The library itself reads none of them.
_ENV_TUNING = {"RTC_BATCHES": "batches", "RTC_BATCH_SAMPLES": "batch_samples", "RTC_PIPES": "pipes", "RTC_BLOCKS_PER_SM": "blocks_per_sm",
               "RTC_NO_TAPER": "no_taper", "RTC_NO_COPY_STREAM": "no_copy_stream", "RTC_NO_STREAM_PRIORITIES": "no_stream_priorities",
               "RTC_NO_GRAPHS": "no_graphs", "RTC_NO_FAST_PLANES": "no_fast_planes", "RTC_NO_PLAIN_SHADE": "no_plain_shade", "RTC_TWO_PASS": "two_pass",
               "RTC_LBVH_LEAF": "lbvh_leaf", "RTC_BVH_BUILDER": "bvh_builder", "RTC_BUILD_THREADS": "build_threads",
               "RTC_LEAF_WEIGHT_X100": "sah_leaf_weight_x100", "RTC_BINS": "sah_bins", "RTC_SMALL_NODE": "sah_small_node"}


def set_tuning(**fields):
    """rtc_set_tuning: process-wide knobs (0 / omitted = library default).  set_tuning() with no argument resets them."""
    lib = load_rtc()
    t = Tuning()
    for k, v in fields.items():
        setattr(t, k, int(v))
    lib.rtc_set_tuning(C.byref(t) if fields else None)


class tuning:
    """with tuning(pipes=3, batch_samples=1024): ...   — knobs set for the block, defaults restored afterwards."""

    def __init__(self, **fields):
        self.fields = fields

    def __enter__(self):
        self.saved = Tuning()
        load_rtc().rtc_get_tuning(C.byref(self.saved))
        t = Tuning.from_buffer_copy(self.saved)
        for k, v in self.fields.items():
            setattr(t, k, int(v))
        load_rtc().rtc_set_tuning(C.byref(t))
        return self

    def __exit__(self, *a):
        load_rtc().rtc_set_tuning(C.byref(self.saved))


STAT_KEYS = ["prims", "bvh_nodes", "bvh_depth", "unbounded", "gates_exact", "gates_conservative", "gates_always_pass",
             "csg_roots", "xforms", "features", "bvh_items", "device_bytes", "cone_guards"]

FLAG_BRUTE_FORCE, FLAG_PROFILE, FLAG_SCATTER_ROWS = 1, 2, 4


def par_render_multi(name, hsize, vsize, n_devices=0, data_dir=None):
    """Image::par_render of the host builder API over n_devices GPUs in ONE process (rtc_render_multi).  Returns (image, Stats)."""
    lib = load_driver()
    out = np.empty((vsize, hsize, 3), dtype=np.float64)
    stats = Stats()
    err = C.create_string_buffer(1024)
    rc = lib.drv_par_render_multi(name.encode(), (data_dir or DATA_DIR).encode(), hsize, vsize, n_devices, out.ctypes.data, C.byref(stats), err, 1024)
    if rc != 0:
        raise RtcError(err.value.decode())
    return out, stats


def lib_paths():
    return {"librtc": os.path.join(LIB_DIR, "librtc.so"), "driver": os.path.join(LIB_DIR, "librtc_driver.so"),
            "oracle": os.path.join(ROOT, "build", "liboracle.so")}


_rtc = None
_drv = None


def load_rtc():
    """Loads the product library; raises if it has not been built (there is no fallback path)."""
    global _rtc
    if _rtc is None:
        path = lib_paths()["librtc"]
        if not os.path.exists(path):
            raise RtcError(f"{path} is missing: run `python -m raytracer_challenge_b200._build` (no CPU fallback exists)")
        lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
        lib.rtc_last_error.restype = C.c_char_p
        lib.rtc_device_count.restype = C.c_int
        lib.rtc_shard_pixels.restype = C.c_uint64
        lib.rtc_shard_pixels.argtypes = [C.c_int32] * 6 + [C.c_void_p]
        lib.rtc_quantize.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        lib.rtc_scene_stat.restype = C.c_int64
        lib.rtc_scene_stat.argtypes = [C.c_void_p, C.c_int]
        lib.rtc_scene_validate.argtypes = [C.c_void_p, C.c_void_p]
        lib.rtc_scene_create.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.rtc_scene_destroy.argtypes = [C.c_void_p]
        lib.rtc_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.rtc_render_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.rtc_render_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.rtc_alloc_host.restype = C.c_void_p
        lib.rtc_alloc_host.argtypes = [C.c_uint64]
        lib.rtc_free_host.argtypes = [C.c_void_p]
        lib.rtc_host_register.argtypes = [C.c_void_p, C.c_uint64]
        lib.rtc_host_unregister.argtypes = [C.c_void_p]
        lib.rtc_microbench.argtypes = [C.c_int, C.c_void_p]
        lib.rtc_set_tuning.argtypes = [C.c_void_p]
        lib.rtc_get_tuning.argtypes = [C.c_void_p]
        _rtc = lib
        env = {field: int(os.environ[name]) for name, field in _ENV_TUNING.items() if os.environ.get(name)}
        if env:
            set_tuning(**env)
    return _rtc


def pinned_array(shape, dtype=np.float64):
    """numpy array backed by page-locked host memory (rtc_alloc_host); keeps its allocation alive."""
    lib = load_rtc()
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    ptr = lib.rtc_alloc_host(n)
    if not ptr:
        raise RtcError(lib.rtc_last_error().decode())
    buf = (C.c_char * n).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
    return arr, ptr


def free_pinned(ptr):
    load_rtc().rtc_free_host(ptr)


def microbench():
    """Measured FP64/FP32 issue rates (TFLOP/s) and L2 read bandwidth (GB/s) on the current device."""
    lib = load_rtc()
    out = {}
    for key, what in (("fp64_fma_tflops", 0), ("fp64_mul_add_tflops", 1), ("fp32_fma_tflops", 2), ("fp32_mul_add_tflops", 3), ("l2_read_gbs", 4)):
        v = C.c_double()
        if lib.rtc_microbench(what, C.byref(v)) != 0:
            raise RtcError("rtc_microbench failed")
        out[key] = v.value
    return out


def load_driver():
    global _drv
    if _drv is None:
        load_rtc()
        path = lib_paths()["driver"]
        if not os.path.exists(path):
            raise RtcError(f"{path} is missing: run `python -m raytracer_challenge_b200._build`")
        lib = C.CDLL(path)
        lib.drv_open.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        lib.drv_close.argtypes = [C.c_void_p]
        lib.drv_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        lib.drv_render_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        lib.drv_par_render_ppm.restype = C.c_int64
        lib.drv_par_render_ppm.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_char_p, C.c_int]
        lib.drv_par_render_multi.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        lib.drv_e2e_step.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        lib.drv_par_render_step.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        lib.drv_scene.restype = C.c_void_p
        lib.drv_scene.argtypes = [C.c_void_p]
        lib.drv_camera.restype = C.c_void_p
        lib.drv_camera.argtypes = [C.c_void_p]
        _drv = lib
    return _drv


def device_count():
    return load_rtc().rtc_device_count()


def quantize(rgb):
    """Color::clamp (src/color.rs:42-46) on an array of f64 -> uint8 of the same shape."""
    a = np.ascontiguousarray(rgb, dtype=np.float64)
    out = np.empty(a.shape, dtype=np.uint8)
    load_rtc().rtc_quantize(a.ctypes.data, out.ctypes.data, a.size)
    return out


def shard_pixels(w, h, shard_index, shard_count, tile=(0, 0)):
    lib = load_rtc()
    n = lib.rtc_shard_pixels(w, h, tile[0], tile[1], shard_index, shard_count, None)
    idx = np.empty(n, dtype=np.uint32)
    lib.rtc_shard_pixels(w, h, tile[0], tile[1], shard_index, shard_count, idx.ctypes.data)
    return idx


class Session:
    """A named scene (scenes/scenes.hpp) built through the host builder API and, if device >= 0, uploaded."""

    def __init__(self, name, hsize, vsize, device=0, data_dir=DATA_DIR):
        self.lib = load_driver()
        self.name, self.hsize, self.vsize, self.device = name, hsize, vsize, device
        self.handle = C.c_void_p()
        stats = (C.c_int64 * 16)()
        err = C.create_string_buffer(1024)
        rc = self.lib.drv_open(name.encode(), data_dir.encode(), hsize, vsize, device, C.byref(self.handle), stats, err, 1024)
        if rc != 0:
            raise RtcError(err.value.decode())
        self.build_stats = {k: int(stats[i]) for i, k in enumerate(STAT_KEYS)}

    def close(self):
        if self.handle:
            self.lib.drv_close(self.handle)
            self.handle = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def out_shape(self, opts):
        w, h = (opts.win_w, opts.win_h) if opts.win_w > 0 and opts.win_h > 0 else (self.hsize, self.vsize)
        if opts.shard_count <= 1:
            return (h, w, 3)
        n = load_rtc().rtc_shard_pixels(w, h, opts.tile_w, opts.tile_h, opts.shard_index, opts.shard_count, None)
        return (n, 3)

    def render(self, opts=None, out=None):
        """rtc_render: host buffer in, host pixels out (f64 rgb).  Returns (image, Stats)."""
        opts = opts or Opts()
        shape = self.out_shape(opts)
        if out is None:
            out = np.empty(shape, dtype=np.float64)
        stats = Stats()
        err = C.create_string_buffer(1024)
        rc = self.lib.drv_render(self.handle, C.byref(opts), out.ctypes.data, C.byref(stats), err, 1024)
        if rc != 0:
            raise RtcError(err.value.decode())
        return out, stats

    def render_device(self, opts=None):
        """rtc_render_device: result stays in HBM.  Returns (device pointer, pixel count, Stats)."""
        opts = opts or Opts()
        dev = C.c_void_p()
        n = C.c_uint64()
        stats = Stats()
        err = C.create_string_buffer(1024)
        rc = self.lib.drv_render_device(self.handle, C.byref(opts), C.byref(dev), C.byref(n), C.byref(stats), err, 1024)
        if rc != 0:
            raise RtcError(err.value.decode())
        return dev.value, n.value, stats
