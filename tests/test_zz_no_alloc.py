"""rtc_render_noalloc (include/rtc.h): the render of a small window with every byte of scratch in a static device array —
no cudaMalloc in the call.  No counterpart in the reference.
Collected last (the file name), so that a failure here cannot hide another test under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

import raytracer_challenge_b200 as R

pytestmark = pytest.mark.gpu


def noalloc_render(sess, opts, shape):
    lib = R.load_rtc()
    out = np.full(shape, np.nan, dtype=np.float64)
    st = R.Stats()
    rc = lib.rtc_render_noalloc(sess.lib.drv_scene(sess.handle), sess.lib.drv_camera(sess.handle), C.byref(opts), out.ctypes.data, C.byref(st))
    if rc != 0:
        raise R.RtcError(lib.rtc_last_error().decode())
    return out, st


def test_two_calls_in_a_row_equal_rtc_render(built):
    """Two calls back to back on different scenes and windows (the second finds the first one's leftovers in the shared array):
    each must equal what the existing entry point, rtc_render, returns for the same window.  Same kernels, same order of
    operations, one writer per colour: the tolerance of the other tests for "same pixels however produced" is exact equality."""
    cases = [("teapot_low", 160, 90, R.Opts(window=(40, 30, 64, 33))),         # reflective mesh: a chain of single children
             ("glass_air_bubble", 64, 64, R.Opts(window=(20, 20, 24, 24))),    # reflect + refract: the binary path tree (fan 2)
             ("csg_noise", 96, 54, R.Opts(window=(8, 6, 16, 12), spp=(2, 2)))]  # CSG, procedural patterns, supersampling, two bins
    for name, w, h, opts in cases:
        with R.Session(name, w, h, device=0) as s:
            shape = (opts.win_h, opts.win_w, 3)
            got, st = noalloc_render(s, opts, shape)
            want, st_ref = s.render(opts)
            again, _ = noalloc_render(s, opts, shape)
        assert np.isfinite(got).all()
        assert np.array_equal(got, want), (name, float(np.abs(got - want).max()))
        assert np.array_equal(again, want), name
        assert (st.rays, st.surface_hits, st.rays_ref) == (st_ref.rays, st_ref.surface_hits, st_ref.rays_ref)


def test_sizes_beyond_the_static_array_are_rejected(built):
    with R.Session("teapot_low", 1920, 1080, device=0) as s:
        with pytest.raises(R.RtcError, match="RTC_NOALLOC_SCRATCH_BYTES"):
            noalloc_render(s, R.Opts(), (1080, 1920, 3))
        with pytest.raises(R.RtcError, match="sharding"):
            noalloc_render(s, R.Opts(shard=(0, 2), window=(0, 0, 16, 16)), (16, 16, 3))
        ok, _ = noalloc_render(s, R.Opts(window=(900, 500, 128, 128)), (128, 128, 3))  # 16 384 samples: fits
        want, _ = s.render(R.Opts(window=(900, 500, 128, 128)))
    assert np.array_equal(ok, want)
