"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol include/rtc.h declares, its
struct layouts match the Python mirror, and the host half of scene upload (validation, flattening, gates, CSG
programs, BVH) accepts / rejects what it should.  No compute call is made: there is no GPU here."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import raytracer_challenge_b200 as R
from common import ROOT, oracle_quantize

HEADER = os.path.join(ROOT, "include", "rtc.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtc_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    names = declared_functions()
    for must in ["rtc_scene_create", "rtc_scene_create_multi", "rtc_scene_destroy", "rtc_render", "rtc_render_device", "rtc_render_multi", "rtc_last_error",
                 "rtc_device_count", "rtc_shard_pixels", "rtc_quantize", "rtc_scene_validate"]:
        assert must in names


def test_library_exports_every_declared_symbol(built):
    lib = C.CDLL(built["librtc"])
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing
    # and nothing torch-typed crosses the boundary: the library does not link torch / python
    ldd = subprocess.run(["ldd", built["librtc"]], capture_output=True, text=True).stdout
    assert "libtorch" not in ldd and "libpython" not in ldd and "libc10" not in ldd, ldd


def test_struct_layouts_match_a_c_compiler(built, tmp_path):
    """sizeof/offsetof from gcc on include/rtc.h against the ctypes mirrors in binding.py."""
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "rtc.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(rtc_opts), sizeof(rtc_stats), sizeof(rtc_camera), sizeof(rtc_prim), sizeof(rtc_node), sizeof(rtc_material),"
                   "sizeof(rtc_pattern), sizeof(rtc_light), offsetof(rtc_stats, ms_trace));return 0;}\n")
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes[0] == C.sizeof(R.Opts)
    assert sizes[1] == C.sizeof(R.Stats)
    assert sizes[8] == R.Stats.ms_trace.offset
    assert sizes[2] == 4 + 4 + 16 * 8 + 3 * 8
    assert C.sizeof(R.Tuning) == 32 * 4  # rtc_tuning: 18 knobs + 14 reserved int32


def test_build_flags_live_in_the_former_padding_of_the_scene_descriptor(built, tmp_path, lib_types):
    """rtc_scene_desc.build_flags was added where the compiler used to pad (seven int32, then pointers): the size and the
    pointer offsets of ABI version 2 are unchanged, and descriptors that never heard of the field read as 0 = default."""
    src = tmp_path / "desc.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "rtc.h"\nint main(){printf("%zu %zu %zu\\n", sizeof(rtc_scene_desc),'
                   "offsetof(rtc_scene_desc, build_flags), offsetof(rtc_scene_desc, prims));return 0;}\n")
    exe = tmp_path / "desc"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    size, off_flags, off_prims = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    Desc = lib_types[5]
    assert (size, off_flags, off_prims) == (32 + 6 * 8, 28, 32)
    assert (C.sizeof(Desc), Desc.build_flags.offset, Desc.prims.offset) == (size, off_flags, off_prims)


def test_no_cuda_device_is_an_error_not_a_fallback(built):
    """The product path has no CPU fallback: opening a scene on a device fails loudly in this GPU-less container."""
    if R.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(R.RtcError):
        R.Session("default_world", 16, 16, device=0)


@pytest.mark.parametrize("name,items,staged", [("teapot", 6320, True), ("benchmark", 1000, True), ("hexagon", 12, False)])
def test_fast_build_stages_the_items_for_the_device_and_builds_no_host_tree(built, name, items, staged):
    """rtc_scene_desc.build_flags & RTC_BUILD_FAST (here through rtc_tuning.bvh_builder = device): scenes with 64 or more BVH
    items leave the tree to the device (bvh_build.cu), so the host half reports no nodes; tiny scenes keep the host build."""
    with R.tuning(bvh_builder=R.BVH_DEVICE):
        s = R.Session(name, 64, 36, device=-1)
        b = s.build_stats
        s.close()
    assert b["bvh_items"] == items
    assert (b["bvh_nodes"] == 0) == staged
    t = R.Tuning()
    R.load_rtc().rtc_get_tuning(C.byref(t))
    assert t.bvh_builder == 0  # the block restored the defaults
    s = R.Session(name, 64, 36, device=-1)
    assert s.build_stats["bvh_nodes"] >= 1
    s.close()


@pytest.mark.parametrize("name", R.SCENES + ["fuzz:1", "fuzz:2", "fuzz:3", "fuzz:7"])
def test_host_scene_build(built, name):
    """rtc_scene_validate = the host half of rtc_scene_create: flatten, dedupe transforms, classify group boxes, build
    CSG programs and the BVH."""
    s = R.Session(name, 64, 36, device=-1)
    b = s.build_stats
    s.close()
    assert b["prims"] > 0
    assert b["bvh_items"] + b["unbounded"] <= b["prims"] + b["csg_roots"]
    if b["bvh_items"] > 1:
        assert b["bvh_nodes"] >= 1 and b["bvh_depth"] < 48
    if name == "teapot":
        assert b["prims"] == 6323 and b["bvh_items"] == 6320 and b["unbounded"] == 3 and b["xforms"] == 4
    if name == "teapot_low":
        assert b["bvh_items"] == 240
    if name == "benchmark":
        assert b["prims"] == 1001 and b["unbounded"] == 1 and b["gates_conservative"] == 9 and b["gates_exact"] == 0
    if name == "hexagon":
        assert b["prims"] == 12 and b["unbounded"] == 0  # finite-range open cylinders get finite bounds
    if name == "csg_noise":
        assert b["csg_roots"] >= 1 and (b["features"] & 2)
    if name == "glass_air_bubble":
        assert b["features"] & 1
    if name == "chapter14_title":  # 36 finite open cones: BVH items behind per-ray guards, only the backdrop plane is tested for every ray
        assert b["unbounded"] == 1 and b["bvh_items"] == 72 and b["cone_guards"] == 36 and (b["features"] & 4)
    if name == "csg_cuts":  # CSG items enter the BVH under the box of what they can report: a cube cut by a half-space and a half-space
        # AND a sphere are bounded items although one operand is a plane; only the floor is tested for every ray
        assert b["csg_roots"] == 7 and b["unbounded"] == 1 and b["bvh_items"] == 8
    if name == "many_gates":  # every box above a finite cone is an exact gate; the 10 outer boxes that also hold a sphere still are
        assert b["gates_exact"] > 32 and b["cone_guards"] >= 1
    if name == "cone_slope":  # boxes above a finite cone stay exact gates: its single-root crossing may lie outside them
        assert b["cone_guards"] == 1 and b["gates_exact"] == 2 and b["gates_conservative"] == 0


def _desc(lib_types, **over):
    """A minimal valid scene description (one sphere, one material, one plain pattern, one light, root node)."""
    Prim, Node, Mat, Pat, Light, Desc = lib_types
    ident = (C.c_double * 16)(*[1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1])
    prim = Prim()
    prim.kind = 0
    prim.casts_shadow = 1
    for f in ("transform_inv", "transform_inv_tsp", "material_inv"):
        setattr(prim, f, ident)
    inf = float("inf")
    node = Node(kind=3, first_child=0, n_children=1)
    node.bbox_min = (C.c_double * 3)(-inf, -inf, -inf)
    node.bbox_max = (C.c_double * 3)(inf, inf, inf)
    mat = Mat(ambient=0.1, diffuse=0.9, specular=0.9, shininess=200.0, refractive_index=1.0, pattern=0)
    pat = Pat(kind=1, left=-1, right=-1)
    pat.transform_inv = ident
    light = Light()
    children = (C.c_int32 * 1)(0)
    keep = [prim, node, mat, pat, light, children]
    d = Desc(abi_version=2, n_prims=1, n_nodes=1, n_children=1, n_materials=1, n_patterns=1, n_lights=1,
             prims=C.pointer(prim), nodes=C.pointer(node), children=children, materials=C.pointer(mat), patterns=C.pointer(pat),
             lights=C.pointer(light))
    for k, v in over.items():
        setattr(d, k, v)
    return d, keep


@pytest.fixture(scope="module")
def lib_types():
    class Prim(C.Structure):
        _fields_ = [("kind", C.c_int32), ("material", C.c_int32), ("casts_shadow", C.c_int32), ("closed", C.c_int32),
                    ("transform_inv", C.c_double * 16), ("transform_inv_tsp", C.c_double * 16), ("material_inv", C.c_double * 16),
                    ("min", C.c_double), ("max", C.c_double), ("p1", C.c_double * 3), ("e1", C.c_double * 3), ("e2", C.c_double * 3),
                    ("n1", C.c_double * 3), ("n2", C.c_double * 3), ("n3", C.c_double * 3)]

    class Node(C.Structure):
        _fields_ = [("kind", C.c_int32), ("first_child", C.c_int32), ("n_children", C.c_int32), ("reserved", C.c_int32),
                    ("bbox_min", C.c_double * 3), ("bbox_max", C.c_double * 3)]

    class Mat(C.Structure):
        _fields_ = [(n, C.c_double) for n in ("ambient", "diffuse", "specular", "shininess", "reflective", "transparency", "refractive_index")] + \
                   [("pattern", C.c_int32), ("reserved", C.c_int32)]

    class Pat(C.Structure):
        _fields_ = [("kind", C.c_int32), ("left", C.c_int32), ("right", C.c_int32), ("noise_kind", C.c_int32), ("octaves", C.c_int32),
                    ("reserved", C.c_int32), ("noise_scale", C.c_double), ("color", C.c_double * 3), ("transform_inv", C.c_double * 16)]

    class Light(C.Structure):
        _fields_ = [("intensity", C.c_double * 3), ("origin", C.c_double * 3)]

    class Desc(C.Structure):
        _fields_ = [("abi_version", C.c_int32), ("n_prims", C.c_int32), ("n_nodes", C.c_int32), ("n_children", C.c_int32),
                    ("n_materials", C.c_int32), ("n_patterns", C.c_int32), ("n_lights", C.c_int32), ("build_flags", C.c_uint32),
                    ("prims", C.POINTER(Prim)), ("nodes", C.POINTER(Node)), ("children", C.POINTER(C.c_int32)),
                    ("materials", C.POINTER(Mat)), ("patterns", C.POINTER(Pat)), ("lights", C.POINTER(Light))]

    return Prim, Node, Mat, Pat, Light, Desc


def test_validate_accepts_minimal_scene_and_rejects_broken_ones(built, lib_types):
    lib = R.load_rtc()
    d, keep = _desc(lib_types)
    assert lib.rtc_scene_validate(C.byref(d), None) == 0, lib.rtc_last_error()
    # error convention: non-zero status + message
    d, keep = _desc(lib_types, abi_version=1)
    assert lib.rtc_scene_validate(C.byref(d), None) != 0 and b"abi_version" in lib.rtc_last_error()
    d, keep = _desc(lib_types)
    keep[0].kind = 9
    assert lib.rtc_scene_validate(C.byref(d), None) != 0 and b"unknown kind" in lib.rtc_last_error()
    d, keep = _desc(lib_types)
    keep[0].material = 5
    assert lib.rtc_scene_validate(C.byref(d), None) != 0 and b"material" in lib.rtc_last_error()
    d, keep = _desc(lib_types)
    keep[0].transform_inv[12] = 0.5  # projective row: not something the reference's constructors can produce
    assert lib.rtc_scene_validate(C.byref(d), None) != 0 and b"affine" in lib.rtc_last_error()
    d, keep = _desc(lib_types)
    keep[1].kind = 0  # a CSG root: needs exactly two children (src/shape.rs:80-83) and cannot be the virtual root
    assert lib.rtc_scene_validate(C.byref(d), None) != 0
    d, keep = _desc(lib_types, n_lights=17)
    assert lib.rtc_scene_validate(C.byref(d), None) != 0 and b"lights" in lib.rtc_last_error()
    assert lib.rtc_scene_validate(None, None) != 0


def test_quantize_is_color_clamp(built):
    """src/color.rs:42-46: clamp to [0, 1], x255, round half away from zero; NaN -> 255 because f64::min returns its non-NaN operand."""
    x = np.array([-1.0, 0.0, 0.5 / 255, 0.499 / 255, 1.5 / 255, 0.5, 1.0, 1.9, np.nan, np.inf, -np.inf, 254.5 / 255])
    q = R.quantize(x)
    assert q.tolist() == [0, 0, 1, 0, 2, 128, 255, 255, 255, 255, 0, 255]
    assert np.array_equal(q, oracle_quantize(x))
    r = np.random.default_rng(7).uniform(-0.2, 1.3, 10000)
    assert np.array_equal(R.quantize(r), oracle_quantize(r))


@pytest.mark.parametrize("w,h,tile,n", [(11, 11, (0, 0), 1), (64, 36, (0, 0), 3), (61, 37, (8, 4), 4), (100, 50, (100, 4), 8), (7, 3, (16, 16), 2), (5, 5, (2, 2), 7)])
def test_shards_partition_the_window(built, w, h, tile, n):
    """Every pixel belongs to exactly one shard; padding slots are marked; slot counts are whole tiles."""
    seen = np.zeros(w * h, dtype=np.int32)
    tw, th = (tile[0] or 8), (tile[1] or 4)
    for i in range(n):
        idx = R.shard_pixels(w, h, i, n, tile)
        assert idx.size % (tw * th) == 0
        real = idx[idx != 0xFFFFFFFF]
        seen[real] += 1
    assert (seen == 1).all()
    assert R.shard_pixels(w, h, n, n, tile).size == 0  # shard index out of range -> empty


# ---- the Rust mirror of the header (rust/src/ffi.rs, unbuilt here: no toolchain) must not drift from include/rtc.h ------
C_TO_RUST = {"int32_t": "i32", "uint32_t": "u32", "uint64_t": "u64", "int64_t": "i64", "double": "f64", "int": "c_int", "uint8_t": "u8"}


def _c_structs():
    """struct name -> [(field, rust-type spelling of its C type)] parsed from include/rtc.h."""
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    out = {}
    for body, name in re.findall(r"typedef struct \w+ \{(.*?)\} (\w+);", text, flags=re.S):
        fields = []
        for decl in body.split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            m = re.match(r"(const )?(\w+)\s*(\*?)\s*(.*)", decl)
            const, ctype, ptr, names = m.groups()
            for n in names.split(","):
                n = n.strip()
                star = ptr or ("*" if n.startswith("*") else "")
                n = n.lstrip("* ")
                arr = re.match(r"(\w+)\[(\d+)\]", n)
                base = C_TO_RUST.get(ctype, ctype)
                if star:
                    ty = ("*const " if const else "*mut ") + base
                elif arr:
                    n, ty = arr.group(1), f"[{base}; {arr.group(2)}]"
                else:
                    ty = base
                fields.append((n, ty))
        out[name] = fields
    return out


def _rust_structs():
    text = re.sub(r"//[^\n]*", "", open(os.path.join(ROOT, "rust", "src", "ffi.rs")).read())
    out = {}
    for attrs, name, body in re.findall(r"((?:#\[[^\]]*\]\s*)+)pub struct (\w+) \{(.*?)\n\}", text, flags=re.S):
        assert "repr(C)" in attrs, f"{name} is not #[repr(C)]"
        out[name] = [(f.strip(), " ".join(t.split())) for f, t in re.findall(r"pub (\w+)\s*:\s*([^,]+),", body)]
    return out


def test_rust_ffi_structs_mirror_the_header():
    c, r = _c_structs(), _rust_structs()
    for name in ["rtc_prim", "rtc_node", "rtc_material", "rtc_pattern", "rtc_light", "rtc_scene_desc", "rtc_camera", "rtc_opts", "rtc_stats", "rtc_tuning"]:
        assert name in r, f"rust/src/ffi.rs lacks {name}"
        assert r[name] == c[name], (name, [x for x in zip(r[name], c[name]) if x[0] != x[1]], len(r[name]), len(c[name]))


def test_rust_ffi_functions_and_constants_mirror_the_header():
    header = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    rust = open(os.path.join(ROOT, "rust", "src", "ffi.rs")).read()
    declared = set(declared_functions())
    bound = set(re.findall(r"pub fn (rtc_\w+)\s*\(", rust))
    assert bound <= declared, bound - declared
    # what Image::par_render needs (rust/src/gpu.rs) must be bound
    for must in ["rtc_device_count", "rtc_scene_create_multi", "rtc_scene_destroy", "rtc_render_multi", "rtc_last_error", "rtc_alloc_host", "rtc_free_host"]:
        assert must in bound, must
    # argument counts agree
    for fn in bound:
        c_args = re.search(r"\b" + fn + r"\s*\(([^)]*)\)", header).group(1)
        r_args = re.search(r"pub fn " + fn + r"\s*\(([^)]*)\)", rust).group(1)
        n_c = 0 if c_args.strip() in ("", "void") else c_args.count(",") + 1
        n_r = 0 if not r_args.strip() else r_args.count(",") + 1
        assert n_c == n_r, (fn, c_args, r_args)
    for const in ["RTC_ABI_VERSION", "RTC_BUILD_FAST"]:
        cv = int(re.search(r"#define " + const + r" (\d+)", header).group(1))
        rv = int(re.search(r"pub const " + const + r": \w+ = (\d+);", rust).group(1))
        assert cv == rv, const
    # and build.rs compiles exactly the units the Python build compiles (csrc/UNITS is the one list)
    assert "UNITS" in open(os.path.join(ROOT, "rust", "build.rs")).read()
    from raytracer_challenge_b200 import _build
    names = [n for n, _ in _build.read_units()]
    assert "bvh_build.cu" in names and "render_kernels.cu" in names
    for n in names:
        assert os.path.exists(os.path.join(ROOT, "raytracer_challenge_b200", "csrc", n)), n


def test_staged_item_boxes_round_outwards(built):
    """The device-build staging rounds the padded item boxes outwards with the double -> float conversion itself (MXCSR
    rounding control, scene_build.cpp: convert_directed) and checks that against the scalar f32_down / f32_up once per
    process; a build whose compiler broke the bracket would fall back to the scalar form and say so in the RTC_TRACE line."""
    import os
    import subprocess
    import sys

    code = ("import raytracer_challenge_b200 as R\n"
            "R.set_tuning(bvh_builder=R.BVH_DEVICE)\n"
            "s = R.Session('teapot', 64, 36, device=-1); print(s.build_stats['bvh_items']); s.close()\n")
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, RTC_TRACE="1", PYTHONPATH=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))),
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "6320" in r.stdout
    assert "device-build staging" in r.stderr and "scalar rounding" not in r.stderr
