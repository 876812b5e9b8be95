"""GPU parity: the CUDA path through the C ABI vs the oracle on the same scenes (pytest -m gpu)."""
import numpy as np
import pytest

import raytracer_challenge_b200 as R
from common import compare_images, oracle_render

pytestmark = pytest.mark.gpu

CASES = [
    ("default_world", 11, 11),          # the reference's own end-to-end test (src/image.rs:128-146)
    ("default_world", 160, 90),
    ("three_spheres", 160, 90),
    ("glass_air_bubble", 128, 128),
    ("hexagon", 192, 100),
    ("benchmark", 192, 100),
    ("teapot_low", 96, 50),
    ("csg_noise", 160, 90),
    # the remaining example binaries
    ("chapter11_title", 192, 100),      # six planes, two glass spheres
    ("chapter12_title", 192, 100),      # cubes only: a room, a glass cube, a mirror
    ("chapter13_title", 192, 100),      # open / closed / glass cylinders
    ("chapter14_title", 192, 100),      # open cones and cylinders in nested groups, four lights
    ("cover", 128, 128),                # glass sphere + 17 cubes in nested groups, two lights
    ("cone_slope", 33, 33),             # rays along a cone's slope: the single-root branch without y-range check (src/shape.rs:811-818)
    # direction components that are exactly zero (walk_inverse): hsize == 1 -> d.x == 0, vsize == 1 -> d.y == 0
    ("axis_aligned", 1, 48),
    ("axis_aligned", 48, 1),
    ("axis_aligned", 1, 1),
    ("axis_aligned", 65, 37),
    ("benchmark_glass", 96, 54),        # the 1 000-sphere scene made of glass (two-pass form: > 512 BVH items): up to ~100 crossings on a line
    ("tri_ghost", 96, 54),              # triangles-only BVH (FEAT_TRIS kernels) with a fan that casts no shadow: nearest-hit shadow queries
    ("csg_cuts", 128, 72),              # CSG items under the box of what they can report; right operands skipped when the left reports nothing
    ("many_gates", 96, 54),             # 80 group boxes over finite cones: 64-bit gate masks, slots assigned after the subtree walk
    ("no_shadow_flat", 96, 54),         # casts_shadow = false in front of a casting shape (src/world.rs:40-47); flat triangles
] + [(f"fuzz:{s}", 96, 54) for s in list(range(1, 13)) + [20, 38, 57, 80, 168, 192]]  # (20 ...: CSG groups with an OPEN left operand — tools/fuzz_sweep.py)


@pytest.mark.parametrize("name,w,h", CASES)
def test_scene_matches_oracle(built, name, w, h):
    ref, counts, _ = oracle_render(name, w, h)
    with R.Session(name, w, h, device=0) as s:
        img, st = s.render(R.Opts())
        brute, _ = s.render(R.Opts(flags=1))
    rep = compare_images(img, ref)
    assert rep["ok"], rep
    # tolerance above is the contract; in practice the f64 device path agrees to rounding noise
    assert rep["max_abs_f64"] < 1e-9, rep
    # BVH culling must not change anything relative to testing every primitive
    assert np.array_equal(img, brute)
    # R_ref: the GPU's count of reference-equivalent World::intersect queries equals the oracle's own count
    assert st.rays_ref == int(counts.sum()), (st.as_dict(), counts)


def test_kernel_variants_chosen_per_scene(built):
    """rtc_scene_stat(FEATURES) of an uploaded scene names the kernel variant: 32 = FEAT_TRIS (every BVH item a triangle, plain
    materials/groups), 16 = FEAT_ONEPASS (transparent, small BVH), 1 / 2 / 4 = transparent / CSG / exact gates."""
    want = {"teapot": 32, "teapot_low": 32, "tri_ghost": 32, "benchmark": 0, "glass_air_bubble": 1 | 16, "benchmark_glass": 1, "many_gates": 4}
    for name, feat in want.items():
        with R.Session(name, 32, 18, device=0) as s:
            assert s.build_stats["features"] == feat, (name, s.build_stats["features"])
    with R.Session("csg_noise", 32, 18, device=0) as s:
        assert s.build_stats["features"] & 2 and s.build_stats["features"] & 16


def test_known_answer_pixel(built):
    with R.Session("default_world", 11, 11, device=0) as s:
        img, _ = s.render()
    assert np.allclose(img[5, 5], [0.38066, 0.47583, 0.28550], atol=1e-5)


# ---- the reference's own rendered images (fixtures: tests/golden/make_golden.py) -------------------------------
import json
import os

from common import ROOT, oracle_quantize

_GOLDEN = os.path.join(ROOT, "tests", "golden")
with open(os.path.join(_GOLDEN, "ref_png_crops.json")) as _f:
    _MANIFEST = json.load(_f)


@pytest.mark.parametrize("entry", _MANIFEST, ids=lambda e: e["key"])
def test_cuda_matches_reference_png_window(built, entry):
    """The bin's scene at the bin's native resolution through the C ABI vs the reference's own output image."""
    crops = np.load(os.path.join(_GOLDEN, "ref_png_crops.npz"))
    want = crops[entry["key"]].astype(np.int32)
    with R.Session(entry["scene"], entry["hsize"], entry["vsize"], device=0) as s:
        img, _ = s.render(R.Opts(window=tuple(entry["window"])))
    d = np.abs(oracle_quantize(img).astype(np.int32) - want).max(axis=-1)
    assert (d <= 1).mean() >= 0.999 and d.max() <= 2, {"max": int(d.max()), "n_diff": int((d > 0).sum())}


# ---- sharding, windows, supersampling, profiling: same pixels whichever way they are produced -------------------
def same_pixels(name, a, b):
    """Bit-identical for every scene: each colour of the path tree has exactly one writer (render_kernels.cu: a ray's
    colour is stored where its parent looks for it, no atomic accumulation), so neither sharding, batching, windows nor
    the order in which warps run can change a bit."""
    return np.array_equal(a, b)


@pytest.mark.parametrize("name,w,h,world,tile", [("glass_air_bubble", 61, 37, 3, (0, 0)), ("teapot_low", 64, 36, 2, (64, 4)),
                                                 ("csg_noise", 50, 30, 4, (50, 4)), ("benchmark", 33, 19, 8, (8, 4))])
def test_shards_reassemble_to_the_whole_frame(built, name, w, h, world, tile):
    with R.Session(name, w, h, device=0) as s:
        whole, _ = s.render(R.Opts())
        frame = np.full((h * w, 3), np.nan)
        for r in range(world):
            part, st = s.render(R.Opts(shard=(r, world), tile=tile))
            idx = R.shard_pixels(w, h, r, world, tile)
            real = idx != 0xFFFFFFFF
            frame[idx[real]] = part[real]
            assert st.rays_primary == int(real.sum())
    assert same_pixels(name, frame.reshape(h, w, 3), whole)


@pytest.mark.parametrize("name,w,h,world", [("teapot_low", 64, 38, 2), ("glass_air_bubble", 48, 27, 3), ("three_spheres", 40, 10, 8)])
def test_scatter_rows_fills_one_shared_frame(built, name, w, h, world):
    """RTC_FLAG_SCATTER_ROWS: every shard writes its bands of rows at their row-major place of ONE host image."""
    from raytracer_challenge_b200 import sharding

    with R.Session(name, w, h, device=0) as s:
        whole, _ = s.render(R.Opts())
        frame = np.full((h, w, 3), np.nan)
        for r in range(world):
            s.render(sharding.band_opts(w, r, world), out=frame)
    assert same_pixels(name, frame, whole)


def test_par_render_multi_one_process(built):
    """Image::par_render of the host builder API (rtc_render_multi underneath) over every visible device."""
    w, h = 96, 54
    ref, _, _ = oracle_render("teapot_low", w, h)
    img, st = R.par_render_multi("teapot_low", w, h, n_devices=0)
    assert compare_images(img, ref)["ok"]
    assert st.rays_primary == w * h
    with R.Session("teapot_low", w, h, device=0) as s:
        one, _ = s.render(R.Opts())
    assert np.array_equal(img, one)


def test_par_render_multi_full_teapot_uses_the_device_build_on_every_gpu(built):
    """The 6 320-triangle mesh needs the 64 KB shared-memory sort of the device-side BVH build (an opt-in attribute that is
    per device): Image::par_render over every visible device must equal a one-device render from a host-built scene."""
    w, h = 128, 72
    img, st = R.par_render_multi("teapot", w, h, n_devices=0)
    with R.Session("teapot", w, h, device=0) as s:
        one, st1 = s.render(R.Opts())
    assert np.array_equal(img, one)
    assert st.rays == st1.rays


def test_window_equals_crop(built):
    with R.Session("glass_air_bubble", 80, 60, device=0) as s:
        whole, _ = s.render(R.Opts())
        win, _ = s.render(R.Opts(window=(13, 7, 31, 22)))
    assert same_pixels("glass_air_bubble", whole[7:29, 13:44], win)


@pytest.mark.parametrize("name,spp", [("three_spheres", (2, 2)), ("csg_noise", (4, 4)), ("glass_air_bubble", (3, 2))])
def test_supersampling_matches_oracle(built, name, spp):
    w, h = 48, 27
    ref, counts, _ = oracle_render(name, w, h, spp=spp)
    with R.Session(name, w, h, device=0) as s:
        img, st = s.render(R.Opts(spp=spp))
    rep = compare_images(img, ref)
    assert rep["ok"] and rep["max_abs_f64"] < 1e-9, rep
    assert st.rays_primary == w * h * spp[0] * spp[1]
    assert st.rays_ref == int(counts.sum())


@pytest.mark.parametrize("fuel", [0, 1, 3])
def test_fuel_matches_oracle(built, fuel):
    ref, counts, _ = oracle_render("glass_air_bubble", 64, 64, fuel=fuel)
    with R.Session("glass_air_bubble", 64, 64, device=0) as s:
        img, st = s.render(R.Opts(fuel=fuel))
    assert compare_images(img, ref)["ok"]
    assert st.rays_ref == int(counts.sum())


def test_profile_flag_and_work_tallies(built):
    with R.Session("teapot_low", 128, 72, device=0) as s:
        a, sa = s.render(R.Opts())
        b, sb = s.render(R.Opts(flags=R.FLAG_PROFILE))
    assert np.array_equal(a, b)
    assert sb.ms_trace > 0 and sb.ms_shadow > 0 and sb.ms_shade > 0
    assert sa.ms_trace == 0
    for st in (sa, sb):
        assert st.surface_hits > 0
        assert st.rays_shadow == st.surface_hits * 2  # two lights (src/bin/chapter15_teapot.rs)
    # the work tallies are only counted under RTC_FLAG_PROFILE (rtc.h)
    assert sb.bvh_node_visits > 0 and sb.triangle_tests > 0 and sb.ray_transforms > 0 and sb.max_node_visits > 0
    assert (sa.bvh_node_visits, sa.triangle_tests, sa.ray_transforms, sa.max_node_visits) == (0, 0, 0, 0)
    assert (sa.rays, sa.rays_reflected, sa.rays_refracted, sa.surface_hits) == (sb.rays, sb.rays_reflected, sb.rays_refracted, sb.surface_hits)


def test_repeated_renders_are_identical_and_batches_do_not_matter(built):
    with R.Session("teapot_low", 160, 90, device=0) as s:
        a, _ = s.render(R.Opts())
        with R.tuning(batch_samples=1024, pipes=3):
            b, st = s.render(R.Opts())
            assert st.kernel_launches > 19 * 4
            with R.tuning(no_graphs=1):  # launch by launch instead of the captured graph
                c, _ = s.render(R.Opts())
        d, _ = s.render(R.Opts())
        # copies issued by the calling thread once a batch's event has fired (rtc_tuning.host_paced_copies), into page-locked and
        # into pageable memory, whole frame and a shard whose bands are scattered to their rows
        frame, _ptr = R.pinned_array((90, 160, 3))
        frame[:] = np.nan
        with R.tuning(host_paced_copies=1, batches=7):
            e, st_e = s.render(R.Opts(), out=frame)
            g, _ = s.render(R.Opts())
            from raytracer_challenge_b200 import sharding
            full = np.full((90, 160, 3), np.nan)
            for r in range(3):
                s.render(sharding.band_opts(160, r, 3), out=full)
        assert st_e.kernel_launches > 0
    assert np.array_equal(a, b) and np.array_equal(a, c) and np.array_equal(a, d)
    assert np.array_equal(a, e) and np.array_equal(a, g) and np.array_equal(a, full)


def test_ppm_through_the_host_api(built):
    """Image::par_render + Image::ppm (src/image.rs:93-112): header, <= 5 pixels per text line, rows restart lines."""
    import ctypes as C

    lib = R.load_driver()
    w, h = 11, 11
    cap = 1 << 16
    buf = C.create_string_buffer(cap)
    err = C.create_string_buffer(1024)
    n = lib.drv_par_render_ppm(b"default_world", R.DATA_DIR.encode(), w, h, 0, buf, cap, err, 1024)
    assert n > 0, err.value
    buf2 = C.create_string_buffer(cap)  # device < 0: through Camera::render(&World), the alias north_star names, over every visible GPU
    assert lib.drv_par_render_ppm(b"default_world", R.DATA_DIR.encode(), w, h, -1, buf2, cap, err, 1024) == n and buf2.raw[:n] == buf.raw[:n]
    text = buf.raw[:n].decode()
    lines = text.split("\n")
    assert lines[0] == "P3" and lines[1] == "11 11" and lines[2] == "255" and text.endswith("\n")
    body = lines[3:-1]
    assert len(body) == 11 * 3  # 11 pixels per row -> 5 + 5 + 1
    assert [len(l.split()) for l in body[:3]] == [15, 15, 3]
    vals = np.array(" ".join(body).split(), dtype=np.int64).reshape(h, w, 3)
    ref, _, _ = oracle_render("default_world", w, h)
    assert np.array_equal(vals, oracle_quantize(ref))


# ---- BASELINE.json sizes: oracle-checked windows inside the full-size frame of every config -------------------------
# (C1's 11x11 and C4's 3840x2160 are covered by test_known_answer_pixel / test_full_size_teapot_properties.)
BASELINE_WINDOWS = [
    # C2: chapter 11 glass air bubble, 1920x1080, fuel 5
    ("glass_air_bubble", 1920, 1080, (1, 1), [(952, 536, 16, 8), (700, 420, 16, 8), (1100, 300, 16, 8)]),
    # C3: chapter 14 hexagon (silhouette edge, jittered stripes) and benchmark (sphere block, backdrop), 3840x2160
    ("hexagon", 3840, 2160, (1, 1), [(2160, 540, 16, 8), (2640, 945, 16, 8), (1200, 675, 16, 8)]),
    ("benchmark", 3840, 2160, (1, 1), [(1900, 1000, 16, 8), (1560, 520, 16, 8)]),
    # C5: CSG + Simplex / fractal noise, 7680x4320 with 4x4 supersampling: a CSG edge, the glass lens (edge, inside), the
    # fractal-noise floor, the cone-union-sphere tower
    ("csg_noise", 7680, 4320, (4, 4), [(1600, 1120, 16, 8), (4160, 1350, 16, 8), (4480, 1350, 16, 8), (3840, 3840, 16, 8), (6000, 1280, 16, 8)]),
]


@pytest.mark.parametrize("name,w,h,spp,windows", BASELINE_WINDOWS, ids=[c[0] for c in BASELINE_WINDOWS])
def test_windows_at_the_baseline_size_match_the_oracle(built, name, w, h, spp, windows):
    with R.Session(name, w, h, device=0) as s:
        for win in windows:
            ref, counts, _ = oracle_render(name, w, h, window=win, spp=spp)
            img, st = s.render(R.Opts(window=win, spp=spp))
            rep = compare_images(img, ref)
            assert rep["ok"] and rep["max_abs_f64"] < 1e-9, (win, rep)
            assert st.rays_ref == int(counts.sum()), (win, st.as_dict(), counts)


def test_full_size_config5_properties(built):
    """7680x4320 x (4x4) samples per pixel = 531 M primary rays (BASELINE.json configs[4]): the frame assembled from 8 band
    shards equals the unsharded frame; windows of the full frame equal the same windows rendered alone (which
    test_windows_at_the_baseline_size_match_the_oracle ties to the oracle); ray-count identities."""
    from raytracer_challenge_b200 import sharding

    w, h, spp = 7680, 4320, (4, 4)
    with R.Session("csg_noise", w, h, device=0) as s:
        whole, st = s.render(R.Opts(spp=spp))
        assert st.rays_primary == w * h * 16
        assert st.rays_shadow == 2 * st.surface_hits  # two lights
        assert st.rays_ref >= st.rays
        assert np.isfinite(whole).all()
        for win in [(1600, 1120, 16, 8), (4480, 1350, 16, 8), (7664, 4312, 16, 8), (0, 0, 16, 8)]:
            alone, _ = s.render(R.Opts(window=win, spp=spp))
            x, y, ww, wh = win
            assert same_pixels("csg_noise", whole[y:y + wh, x:x + ww], alone), win
        frame = np.full((h, w, 3), np.nan)
        for r in range(8):
            s.render(sharding.band_opts(w, r, 8, spp=spp), out=frame)
        assert same_pixels("csg_noise", frame, whole)


# ---- full-size properties (BASELINE.json sizes; the brute-force oracle cannot finish these in seconds) -----------
def test_full_size_teapot_properties(built):
    """3840x2160 teapot, fuel 5: (a) the frame assembled from 8 band shards equals the unsharded frame bit for bit;
    (b) a window rendered alone equals the crop; (c) an oracle-checked window inside the full frame; (d) ray counts obey
    the identities of SURVEY.md §8(d): shadow = hits * L, R_ref >= R_uniq."""
    from raytracer_challenge_b200 import sharding

    w, h = 3840, 2160
    with R.Session("teapot", w, h, device=0) as s:
        whole, st = s.render(R.Opts())
        frame = np.full((h, w, 3), np.nan)
        for r in range(8):
            s.render(sharding.band_opts(w, r, 8), out=frame)
        assert np.array_equal(frame, whole)
        win, _ = s.render(R.Opts(window=(1900, 1000, 64, 32)))
        assert np.array_equal(whole[1000:1032, 1900:1964], win)
    ref, _, _ = oracle_render("teapot", w, h, window=(1900, 1000, 24, 12))
    rep = compare_images(whole[1000:1012, 1900:1924], ref)
    assert rep["ok"] and rep["max_abs_f64"] < 1e-9, rep
    assert st.rays_primary == w * h
    assert st.rays_shadow == 2 * st.surface_hits
    assert st.rays_ref >= st.rays
    assert np.isfinite(whole).all()


@pytest.mark.parametrize("name", ["teapot_low", "chapter11_title", "glass_air_bubble", "three_spheres"])
def test_planes_in_kernel_parameters_equal_the_general_list(built, name):
    """Scenes whose always-tested list is ungated planes keep them in the kernel parameters (device_scene.h,
    FastPlane); rtc_tuning.no_fast_planes forces the general unbounded[] -> prims[] -> xforms[] walk.  Same expressions, so the
    frames must be bit-identical (up to the order of the atomic adds where one pixel has several nodes in a level)."""
    w, h = 96, 54
    with R.Session(name, w, h, device=0) as s:
        fast, st = s.render(R.Opts())
    with R.tuning(no_fast_planes=1), R.Session(name, w, h, device=0) as s:
        general, st2 = s.render(R.Opts())
    assert same_pixels(name, fast, general)
    assert (st.rays, st.surface_hits) == (st2.rays, st2.surface_hits)


@pytest.mark.parametrize("name,w,h", [("teapot_low", 96, 54), ("benchmark", 96, 54), ("chapter14_title", 96, 54), ("teapot", 160, 90), ("fuzz:3", 96, 54)])
def test_device_built_bvh_renders_the_same_frame(built, name, w, h):
    """RTC_BUILD_FAST (what Image::par_render passes: the scene lives for one frame) builds the BVH on the device
    (bvh_build.cu, LBVH) instead of the host's binned-SAH build.  A BVH only culls: the frame must equal the SAH
    build's and the brute-force scan's bit for bit."""
    with R.Session(name, w, h, device=0) as s:
        sah, st = s.render(R.Opts())
        host_stats = dict(s.build_stats)
    with R.tuning(bvh_builder=R.BVH_DEVICE), R.Session(name, w, h, device=0) as s:
        lbvh, st2 = s.render(R.Opts())
        brute, _ = s.render(R.Opts(flags=1))
        dev_stats = dict(s.build_stats)
    assert same_pixels(name, sah, lbvh) and same_pixels(name, lbvh, brute)
    assert (st.rays, st.surface_hits) == (st2.rays, st2.surface_hits)
    if host_stats["bvh_items"] >= 64:  # smaller scenes keep the host build
        assert dev_stats["bvh_depth"] >= 3 and 0 < dev_stats["bvh_nodes"] < dev_stats["bvh_items"]
        assert dev_stats["device_bytes"] > host_stats["device_bytes"]  # the arena also holds the builder's scratch: the device path ran


# ---- capacity: scenes beyond the sizes the example binaries need ---------------------------------------------------------
def test_hundred_thousand_triangle_mesh_on_the_device_built_bvh(built):
    """src/obj.rs has no size limit.  teapot_high tessellated to 101 120 smooth triangles: the device-side build sorts its
    Morton keys with the multi-block radix sort (bvh_build.cu) instead of the one-block sort; windows on the mesh equal
    the oracle, the host-built SAH tree and the brute-force scan."""
    w, h = 3840, 2160
    windows = [(1900, 1300, 12, 6), (1700, 1400, 12, 6)]
    with R.Session("teapot_tess", w, h, device=0) as s:
        host_stats = dict(s.build_stats)
        sah = [s.render(R.Opts(window=win))[0] for win in windows]
    with R.tuning(bvh_builder=R.BVH_DEVICE), R.Session("teapot_tess", w, h, device=0) as s:
        dev_stats = dict(s.build_stats)
        for win, want in zip(windows, sah):
            img, st = s.render(R.Opts(window=win))
            brute, _ = s.render(R.Opts(window=win, flags=1))
            ref, counts, _ = oracle_render("teapot_tess", w, h, window=win)
            rep = compare_images(img, ref)
            assert rep["ok"] and rep["max_abs_f64"] < 1e-9, (win, rep)
            assert np.array_equal(img, want) and np.array_equal(img, brute), win
            assert st.rays_ref == int(counts.sum())
        whole, st = s.render(R.Opts())
        assert np.isfinite(whole).all() and st.rays_primary == w * h
    assert host_stats["bvh_items"] == 101120 and dev_stats["bvh_items"] == 101120
    assert 0 < dev_stats["bvh_nodes"] < 101120 and 3 <= dev_stats["bvh_depth"] <= 62
    assert dev_stats["device_bytes"] > host_stats["device_bytes"]  # the arena also holds the builder's scratch: the device path ran


@pytest.mark.parametrize("two_pass", [0, 1])
def test_replays_that_overflow_the_on_chip_lists_fail_loudly(built, two_pass):
    """glass_shells: the camera sits inside 70 concentric glass shells, so every n1 / n2 replay sees 70+ crossings and up to 70
    open containers — more than the on-chip lists hold (64 / 24, device_scene.h).  That is a capacity limit of this
    implementation, not of the reference ("Capacity"): the call must FAIL with an error in both forms of the walk,
    never return a wrong pixel.  (An in-kernel retry with lists in global scratch was built and measured: the mere presence
    of the fallback code cost the transparent kernels 15-20 %, so it was not kept.)"""
    with R.tuning(two_pass=two_pass), R.Session("glass_shells", 48, 27, device=0) as s:
        assert bool(s.build_stats["features"] & 16) == (not two_pass)  # FEAT_ONEPASS
        with pytest.raises(R.RtcError, match="overflow"):
            s.render(R.Opts())
