"""CPU tests of the steps either side of the render path, through the host API:

* ObjParser: the cases of the reference's own parser tests (src/obj.rs:289-668), same inputs, same expectations;
* Image::ppm: the text format of src/image.rs:93-112, against a line-by-line Python restatement;
* Image::png: the replacement of the bins' `zsh -c pnmtopng` child process, decoded here with zlib.

No GPU needed."""
import ctypes as C
import json
import os
import struct
import zlib

import numpy as np
import pytest

import raytracer_challenge_b200 as R
from common import oracle_quantize


def _drv():
    lib = R.load_driver()
    lib.drv_obj_parse.restype = C.c_int64
    lib.drv_obj_parse.argtypes = [C.c_char_p, C.c_int64, C.c_void_p, C.c_int64, C.c_char_p, C.c_int]
    lib.drv_image_encode.restype = C.c_int64
    lib.drv_image_encode.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64]
    return lib


def parse(text):
    lib = _drv()
    raw = text if isinstance(text, bytes) else text.encode()
    cap = 1 << 22
    buf = C.create_string_buffer(cap)
    err = C.create_string_buffer(1024)
    n = lib.drv_obj_parse(raw, len(raw), buf, cap, err, 1024)
    if n < 0:
        raise RuntimeError(err.value.decode())
    assert n <= cap
    return json.loads(buf.raw[:n].decode())


def encode(rgb, w, h, kind):
    lib = _drv()
    a = np.ascontiguousarray(rgb, dtype=np.float64)
    cap = 64 + 20 * a.size + 2 * h + w * h * 4
    buf = C.create_string_buffer(cap)
    n = lib.drv_image_encode(a.ctypes.data, w, h, kind, buf, cap)
    assert 0 < n <= cap
    return buf.raw[:n]


# ---- ObjParser: the reference's own test cases ---------------------------------------------------------------

def test_obj_ignored_lines(built):
    """src/obj.rs:289-329 — five lines of gibberish are five ignored (line number, text) pairs."""
    lines = ["There was a young lady named Bright", "who traveled much faster than light.", "She set out one day",
             "in a relative way,", "and came back the previous night."]
    d = parse("\n".join(lines) + "\n")
    assert d["ignored"] == [[i + 1, s] for i, s in enumerate(lines)]
    assert d["vertices"] == [] and d["triangles"] == [] and d["element"] is None


def test_obj_vertex_records(built):
    """src/obj.rs:331-369"""
    d = parse("v -1 1 0\nv -1.0000 0.5000 0.0000\nv 1 0 0\nv 1 1 0\n")
    assert d["vertices"] == [[-1, 1, 0, 1], [-1, 0.5, 0, 1], [1, 0, 0, 1], [1, 1, 0, 1]]
    assert d["ignored"] == []


def test_obj_triangle_faces(built):
    """src/obj.rs:371-423 — the blank line is reported as ignored line 5; faces keep 1-based indices without normals."""
    d = parse("v -1 1 0\nv -1 0 0\nv 1 0 0\nv 1 1 0\n\nf 1 2 3\nf 1 3 4\n")
    assert len(d["vertices"]) == 4
    assert d["ignored"] == [[5, ""]]
    assert [(t["p"], t["n"]) for t in d["triangles"]] == [([1, 2, 3], None), ([1, 3, 4], None)]
    assert all(t["group"] == 0 for t in d["triangles"]) and d["groups"] == ["Default"]


def test_obj_triangle_groups(built):
    """src/obj.rs:425-487"""
    d = parse("v -1 1 0\nv -1 0 0\nv 1 0 0\nv 1 1 0\n\ng FirstGroup\nf 1 2 3\ng SecondGroup\nf 1 3 4\n")
    assert d["groups"] == ["Default", "FirstGroup", "SecondGroup"]
    assert [(t["p"], t["group"]) for t in d["triangles"]] == [([1, 2, 3], 1), ([1, 3, 4], 2)]
    assert d["ignored"] == [[5, ""]]


def test_obj_vertex_normal_records(built):
    """src/obj.rs:489-519"""
    d = parse("vn 0 0 1\nvn 0.707 0 -0.707\nvn 1 2 3\n")
    assert d["normals"] == [[0, 0, 1, 0], [0.707, 0, -0.707, 0], [1, 2, 3, 0]]


def test_obj_smooth_triangle_faces(built):
    """src/obj.rs:521-595 — `v//n` and `v/<anything>/n` both yield (vertex, Some(normal)) corners."""
    d = parse("v 0 1 0\nv -1 0 0\nv 1 0 0\n\nvn -1 0 0\nvn 1 0 0\nvn 0 1 0\n\nf 1//3 2//1 3//2\nf 1/0/3 2/102/1 3/14/2\n")
    assert d["ignored"] == [[4, ""], [8, ""]]
    assert [(t["p"], t["n"]) for t in d["triangles"]] == [([1, 2, 3], [3, 1, 2])] * 2
    kinds = [c["kind"] for c in d["element"]["children"]]
    assert kinds == [6, 6]  # RTC_SMOOTH_TRIANGLE


def test_obj_fan_triangulation(built):
    """src/obj.rs:115-123,597-608 — a pentagon becomes (1,2,3), (1,3,4), (1,4,5)."""
    d = parse("v 0 0 0\nv 1 0 0\nv 2 1 0\nv 1 2 0\nv 0 1 0\nf 1 2 3 4 5\n")
    assert [t["p"] for t in d["triangles"]] == [[1, 2, 3], [1, 3, 4], [1, 4, 5]]


def test_obj_element_structure(built):
    """src/obj.rs:610-667 — two named groups: an Aggregation of two Aggregations of one flat triangle each (the
    reference accepts either order because it drains a HashMap); ignored[0] == (5, "")."""
    d = parse("v -1 1 0\nv -1 0 0\nv 1 0 0\nv 1 1 0\n\ng FirstGroup\nf 1 2 3\ng SecondGroup\nf 1 3 4\n")
    assert d["ignored"][0] == [5, ""]
    e = d["element"]
    assert e["group"] == 3 and len(e["children"]) == 2  # RTC_AGGREGATION
    v1, v2, v3, v4 = [-1, 1, 0], [-1, 0, 0], [1, 0, 0], [1, 1, 0]
    got = []
    for g in e["children"]:
        assert g["group"] == 3 and len(g["children"]) == 1 and g["children"][0]["kind"] == 5  # RTC_TRIANGLE
        got.append(g["children"][0]["p"])
    assert got in ([[v1, v2, v3], [v1, v3, v4]], [[v1, v3, v4], [v1, v2, v3]])


def test_obj_single_group_is_not_wrapped(built):
    """src/obj.rs:250-257 — with one non-empty group the group itself is returned."""
    d = parse("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    e = d["element"]
    assert e["group"] == 3 and [c["kind"] for c in e["children"]] == [5]


def test_obj_numbers_and_trailing_text(built):
    """nom's `double` (lexical-core 0.7.4) accepts signs, exponents and bare fractions; text after a complete record
    is ignored because the reference's line parser does not demand end of line (src/obj.rs:53-113).  CRLF files parse
    like LF files."""
    d = parse("v +1.5e0 -2.5E-1 .5 trailing words\r\nv 1e2 0 0\r\nvn 1 0 0\r\nvt 0.5 0.5\r\n")
    assert d["vertices"] == [[1.5, -0.25, 0.5, 1], [100, 0, 0, 1]]
    assert d["normals"] == [[1, 0, 0, 0]]
    assert d["ignored"] == [[4, "vt 0.5 0.5"]]


def test_obj_bad_index_is_an_error(built):
    """The reference indexes `vertices[i - 1]` and panics out of range (src/obj.rs:214-236); here: an exception."""
    with pytest.raises(RuntimeError, match="index out of range"):
        parse("v 0 0 0\nv 1 0 0\nf 1 2 3\n")


def read_mesh(path):
    """RTCMESH1 fixture written by oracle/obj2mesh.cpp from the reference's OBJ files."""
    with open(path, "rb") as f:
        raw = f.read()
    assert raw[:8] == b"RTCMESH1"
    nv, nn, nt, ng = struct.unpack_from("<4I", raw, 8)
    at = 24
    v = np.frombuffer(raw, dtype="<f8", count=3 * nv, offset=at).reshape(nv, 3)
    at += 24 * nv
    n = np.frombuffer(raw, dtype="<f8", count=3 * nn, offset=at).reshape(nn, 3)
    at += 24 * nn
    t = np.frombuffer(raw, dtype="<u4", count=7 * nt, offset=at).reshape(nt, 7)
    return v, n, t, ng


@pytest.mark.parametrize("name", ["teapot_low", "teapot_high"])
def test_obj_parser_reproduces_the_shipped_mesh_fixtures(built, name):
    """data/*.mesh are the reference's obj/teapot_*.obj parsed by the ORACLE's parser (the GPU box has no
    /root/reference).  Where the reference is mounted, the PRODUCT's ObjParser on the same file must give the same
    vertices (bit for bit), normals, fan triangles and groups."""
    ref = f"/root/reference/obj/{name}.obj"
    if not os.path.exists(ref):
        pytest.skip("reference not mounted")
    with open(ref, "rb") as f:
        d = parse(f.read())
    v, n, t, ng = read_mesh(os.path.join(R.DATA_DIR, name + ".mesh"))
    assert np.array_equal(np.array(d["vertices"])[:, :3], v) and np.array_equal(np.array(d["normals"])[:, :3], n)
    got = np.array([tr["p"] + (tr["n"] or [0, 0, 0]) + [tr["group"]] for tr in d["triangles"]], dtype=np.uint32)
    assert np.array_equal(got, t)
    assert len(d["groups"]) == ng


# ---- Image::ppm / Image::png ---------------------------------------------------------------------------------

def ppm_restatement(rgb, w, h):
    """src/image.rs:93-112, line by line."""
    q = oracle_quantize(np.asarray(rgb, dtype=np.float64).reshape(-1)).reshape(-1, 3)
    out = f"P3\n{w} {h}\n255"
    j = 0
    for i, (r, g, b) in enumerate(q):
        if i % w == 0 or j % 5 == 0:
            out += f"\n{r} {g} {b}"
            j = 1
        else:
            out += f" {r} {g} {b}"
            j += 1
    return out + "\n"


@pytest.mark.parametrize("w,h", [(1, 1), (5, 3), (7, 2), (11, 11), (12, 1)])
def test_ppm_matches_the_reference_format(built, w, h):
    rgb = np.random.default_rng(w * 100 + h).uniform(-0.3, 1.4, (h, w, 3))
    assert encode(rgb, w, h, 0).decode() == ppm_restatement(rgb, w, h)


def test_ppm_known_answer(built):
    """The book's canvas example: 5x3, three coloured pixels, channels clamped and scaled to 0..255."""
    rgb = np.zeros((3, 5, 3))
    rgb[0, 0] = [1.5, 0, 0]
    rgb[1, 2] = [0, 0.5, 0]
    rgb[2, 4] = [-0.5, 0, 1]
    assert encode(rgb, 5, 3, 0).decode() == ("P3\n5 3\n255\n255 0 0 0 0 0 0 0 0 0 0 0 0 0 0\n0 0 0 0 0 0 0 128 0 0 0 0 0 0 0\n"
                                             "0 0 0 0 0 0 0 0 0 0 0 0 0 0 255\n")


def decode_png(data):
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    at, chunks = 8, []
    while at < len(data):
        n, tag = struct.unpack(">I4s", data[at:at + 8])
        body = data[at + 8:at + 8 + n]
        (crc,) = struct.unpack(">I", data[at + 8 + n:at + 12 + n])
        assert crc == zlib.crc32(tag + body)
        chunks.append((tag, body))
        at += 12 + n
    assert [t for t, _ in chunks][0] == b"IHDR" and chunks[-1] == (b"IEND", b"")
    w, h, depth, ctype, comp, flt, interlace = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, ctype, comp, flt, interlace) == (8, 2, 0, 0, 0)
    raw = zlib.decompress(b"".join(b for t, b in chunks if t == b"IDAT"))
    rows = np.frombuffer(raw, dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert (rows[:, 0] == 0).all()
    return rows[:, 1:].reshape(h, w, 3)


@pytest.mark.parametrize("w,h", [(1, 1), (5, 3), (64, 36), (300, 200)])
def test_png_decodes_to_the_clamped_pixels(built, w, h):
    """300x200x3 + filter bytes > 65 535 bytes: exercises the multi-block stored-deflate path."""
    rgb = np.random.default_rng(w + h).uniform(-0.3, 1.4, (h, w, 3))
    img = decode_png(encode(rgb, w, h, 1))
    assert np.array_equal(img, oracle_quantize(rgb.reshape(-1)).reshape(h, w, 3))
