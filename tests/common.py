"""Shared test helpers: ctypes access to the oracle (test infrastructure) and the parity comparator."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA_DIR = os.path.join(ROOT, "data")
_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(os.path.join(ROOT, "build", "liboracle.so"))
        lib.orc_render.argtypes = [C.c_char_p, C.c_char_p] + [C.c_int] * 10 + [C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        lib.orc_quantize.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        lib.orc_simplex.restype = C.c_double
        lib.orc_simplex.argtypes = [C.c_double] * 3
        lib.orc_fractal.restype = C.c_double
        lib.orc_fractal.argtypes = [C.c_double] * 3 + [C.c_int]
        _oracle = lib
    return _oracle


def oracle_render(name, hsize, vsize, window=None, fuel=5, spp=(1, 1), threads=0):
    """Reference algorithm on the CPU.  Returns (image[h,w,3] f64, counts[4] = primary/shadow/reflected/refracted, seconds)."""
    x0, y0, w, h = window or (0, 0, hsize, vsize)
    out = np.zeros((h, w, 3), dtype=np.float64)
    counts = np.zeros(4, dtype=np.uint64)
    sec = C.c_double()
    err = C.create_string_buffer(1024)
    rc = oracle().orc_render(name.encode(), DATA_DIR.encode(), hsize, vsize, x0, y0, w, h, fuel, spp[0], spp[1], threads,
                             out.ctypes.data, counts.ctypes.data, C.byref(sec), err, 1024)
    if rc != 0:
        raise RuntimeError("oracle: " + err.value.decode())
    return out, counts, sec.value


def oracle_quantize(img):
    a = np.ascontiguousarray(img, dtype=np.float64)
    out = np.empty(a.shape, dtype=np.uint8)
    oracle().orc_quantize(a.ctypes.data, out.ctypes.data, a.size)
    return out


def compare_images(a, b):
    """The tolerance of BASELINE.json's north_star, in Color::clamp units (src/color.rs:42-46).

    Returns a dict with the LSB histogram over pixels (max channel difference) and the differing pixels.
    pass <=> >= 99.9 % of pixels within 1 LSB and all pixels within 2 LSB.
    """
    qa = oracle_quantize(a).astype(np.int32)
    qb = oracle_quantize(b).astype(np.int32)
    d = np.abs(qa - qb).max(axis=-1)
    n = d.size
    hist = {"0": int((d == 0).sum()), "1": int((d == 1).sum()), "2": int((d == 2).sum()), ">2": int((d > 2).sum())}
    ys, xs = np.nonzero(d.reshape(a.shape[:-1]) > 0) if a.ndim == 3 else (np.nonzero(d > 0)[0], None)
    differing = []
    for k in range(min(len(ys), 20)):
        idx = (ys[k], xs[k]) if xs is not None else (ys[k],)
        differing.append({"at": tuple(int(v) for v in idx), "gpu": qa[idx].tolist(), "oracle": qb[idx].tolist()})
    ok = (hist["0"] + hist["1"]) >= 0.999 * n and hist[">2"] == 0
    return {"ok": bool(ok), "pixels": n, "hist": hist, "differing": differing,
            "max_abs_f64": float(np.nanmax(np.abs(np.asarray(a) - np.asarray(b)))) if n else 0.0}
