"""Asset path of the raytracing host: OBJ/MTL -> mesh arrays, PNG -> texels, PPM out
(vortex-raytracing_b200/host/asset_io.cpp through its C ABI).

Checked (1) on files written by this test with independent Python code (zlib for PNG), and (2) against the reference's own
loaders — Mesh(objFile) over tinyobjloader, Surface(file) over stb_image — compiled into oracle/_ref/libref_assets.so, on the
reference's shipped assets (copied next to it by oracle/Makefile; skipped when oracle/_ref was not built)."""
import ctypes as C
import glob
import importlib
import os
import struct
import zlib

import numpy as np
import pytest

from rt_common import ROOT

host = importlib.import_module("vortex-raytracing_b200.host")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_assets.so")
REF_ASSETS = os.path.join(ROOT, "oracle", "_ref", "assets")


# ---------------------------------------------------------------------------------------------------- PNG
def png_bytes(img, ctype, depth=8, palette=None, filters=(0, 1, 2, 3, 4), level=6, split=0):
    """img: (h, w, channels) uint8/uint16 samples (or (h, w) indices / gray).  Applies the given scanline filters in turn."""
    h, w = img.shape[:2]
    ch = 1 if img.ndim == 2 else img.shape[2]
    if depth == 16:
        rows = img.astype(">u2").reshape(h, -1).view(np.uint8)
    elif depth == 8:
        rows = img.astype(np.uint8).reshape(h, -1)
    else:   # pack sub-byte samples, most significant bits first
        per = 8 // depth
        pad = (-w) % per
        a = np.pad(img.astype(np.uint8).reshape(h, w), ((0, 0), (0, pad)))
        rows = np.zeros((h, (w + pad) // per), np.uint8)
        for k in range(per):
            rows |= (a[:, k::per] << (8 - depth * (k + 1))).astype(np.uint8)
    bpp = max(1, ch * depth // 8)
    raw = bytearray()
    prev = np.zeros(rows.shape[1], np.int32)
    for y in range(h):
        cur = rows[y].astype(np.int32)
        ft = filters[y % len(filters)]
        a = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]]) if bpp < len(cur) else np.zeros_like(cur)
        c = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]]) if bpp < len(cur) else np.zeros_like(cur)
        b = prev
        if ft == 0:
            f = cur
        elif ft == 1:
            f = cur - a
        elif ft == 2:
            f = cur - b
        elif ft == 3:
            f = cur - ((a + b) >> 1)
        else:
            p = a + b - c
            pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
            pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, b, c))
            f = cur - pred
        raw.append(ft)
        raw += (f & 0xFF).astype(np.uint8).tobytes()
        prev = cur

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xFFFFFFFF)

    z = zlib.compress(bytes(raw), level)
    out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 0))
    if palette is not None:
        out += chunk(b"PLTE", np.asarray(palette, np.uint8).tobytes())
    if split:   # several IDAT chunks
        for i in range(0, len(z), split):
            out += chunk(b"IDAT", z[i:i + split])
    else:
        out += chunk(b"IDAT", z)
    return out + chunk(b"IEND", b"")


def pack_rgb(rgb):
    rgb = rgb.astype(np.uint32)
    return (rgb[..., 0] << 16) + (rgb[..., 1] << 8) + rgb[..., 2]


@pytest.mark.parametrize("level", [0, 1, 6, 9])
def test_png_rgb_all_filters(tmp_path, level):
    rng = np.random.default_rng(level)
    img = rng.integers(0, 256, size=(37, 53, 3), dtype=np.uint8)
    img[5:20, 10:40] = (200, 30, 30)                      # runs, so that the compressor emits matches
    p = tmp_path / "rgb.png"
    p.write_bytes(png_bytes(img, 2, level=level, split=777 if level == 6 else 0))
    assert np.array_equal(host.load_png(str(p)), pack_rgb(img))


def test_png_other_colour_types(tmp_path):
    rng = np.random.default_rng(7)
    rgba = rng.integers(0, 256, size=(9, 11, 4), dtype=np.uint8)
    (tmp_path / "rgba.png").write_bytes(png_bytes(rgba, 6))
    assert np.array_equal(host.load_png(str(tmp_path / "rgba.png")), pack_rgb(rgba[..., :3]))        # alpha dropped
    gray = rng.integers(0, 256, size=(8, 13), dtype=np.uint8)
    (tmp_path / "gray.png").write_bytes(png_bytes(gray, 0))
    assert np.array_equal(host.load_png(str(tmp_path / "gray.png")), pack_rgb(np.repeat(gray[..., None], 3, axis=2)))
    ga = rng.integers(0, 256, size=(6, 7, 2), dtype=np.uint8)
    (tmp_path / "ga.png").write_bytes(png_bytes(ga, 4))
    assert np.array_equal(host.load_png(str(tmp_path / "ga.png")), pack_rgb(np.repeat(ga[..., :1], 3, axis=2)))
    pal = rng.integers(0, 256, size=(16, 3), dtype=np.uint8)
    for depth in (8, 4):
        idx = rng.integers(0, 16, size=(10, 15), dtype=np.uint8)
        (tmp_path / "pal.png").write_bytes(png_bytes(idx, 3, depth=depth, palette=pal))
        assert np.array_equal(host.load_png(str(tmp_path / "pal.png")), pack_rgb(pal[idx])), depth
    rgb16 = rng.integers(0, 65536, size=(5, 6, 3), dtype=np.uint16)
    (tmp_path / "rgb16.png").write_bytes(png_bytes(rgb16, 2, depth=16))
    assert np.array_equal(host.load_png(str(tmp_path / "rgb16.png")), pack_rgb(rgb16 >> 8))
    bits = rng.integers(0, 2, size=(7, 19), dtype=np.uint8)
    (tmp_path / "g1.png").write_bytes(png_bytes(bits, 0, depth=1))
    assert np.array_equal(host.load_png(str(tmp_path / "g1.png")), pack_rgb(np.repeat((bits * 255)[..., None], 3, axis=2)))


def test_png_errors(tmp_path):
    with pytest.raises(RuntimeError):
        host.load_png(str(tmp_path / "missing.png"))
    (tmp_path / "junk.png").write_bytes(b"not a png at all")
    with pytest.raises(RuntimeError):
        host.load_png(str(tmp_path / "junk.png"))
    good = png_bytes(np.zeros((4, 4, 3), np.uint8), 2)
    (tmp_path / "cut.png").write_bytes(good[:len(good) // 2])
    with pytest.raises(RuntimeError):
        host.load_png(str(tmp_path / "cut.png"))


# ---------------------------------------------------------------------------------------------------- PPM
def test_ppm_writers(tmp_path):
    rng = np.random.default_rng(3)
    frame = rng.integers(0, 1 << 24, size=(5, 7), dtype=np.uint32)
    host.write_ppm(str(tmp_path / "a.ppm"), frame, binary=False)
    # the reference's writer (tracer.cpp:15-33): P3, rows bottom-up, "R G B\n" per pixel
    want = "P3\n7 5\n255\n" + "".join(f"{(p >> 16) & 255} {(p >> 8) & 255} {p & 255}\n" for row in frame[::-1] for p in row)
    assert (tmp_path / "a.ppm").read_text() == want
    host.write_ppm(str(tmp_path / "b.ppm"), frame, binary=True)
    raw = (tmp_path / "b.ppm").read_bytes()
    head = b"P6\n7 5\n255\n"
    assert raw.startswith(head)
    px = np.frombuffer(raw[len(head):], np.uint8).reshape(5, 7, 3)
    assert np.array_equal(pack_rgb(px), frame[::-1])
    host.write_ppm(str(tmp_path / "c.ppm"), frame, binary=True, flip_vertical=False)
    px = np.frombuffer((tmp_path / "c.ppm").read_bytes()[len(head):], np.uint8).reshape(5, 7, 3)
    assert np.array_equal(pack_rgb(px), frame)


# ---------------------------------------------------------------------------------------------------- OBJ / MTL
OBJ = """# comment
mtllib scene.mtl
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0.5 1.5 0
v 0 0 1
vt 0 0
vt 1 0
vt 1 1
vn 0 0 1
vn 0 1 0
usemtl red
f 1/1/1 2/2/1 3/3/1
usemtl tex
f 1//2 2//2 3//2 4//2
usemtl nosuch
f -6 -5 -4
g pentagon
usemtl red
f 1 2 3 5 4
f 1 2
"""
MTL = """newmtl red
Ka 0.1 0.2 0.3
Kd 0.9 0.1 0.1
Ks 0.5 0.5 0.5
Ke 0 0 1
Ns 32
Ni 1.5
d 0.75
illum 2
newmtl tex
map_Kd -s 1 1 1 checker.png
Tr 0.25
"""


def test_obj_mtl_loader(tmp_path):
    (tmp_path / "scene.obj").write_text(OBJ)
    (tmp_path / "scene.mtl").write_text(MTL)
    tex = np.array([[[255, 0, 0], [0, 255, 0]], [[0, 0, 255], [255, 255, 255]]], dtype=np.uint8)
    (tmp_path / "checker.png").write_bytes(png_bytes(tex, 2))
    m = host.load_obj(str(tmp_path / "scene.obj"))
    P = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0.5, 1.5, 0], [0, 0, 1]], dtype=np.float32)
    # triangle; quad 1-2-3-4 (diagonals equal -> [0,1,3],[1,2,3]); triangle by negative indices; pentagon ear-clipped
    # (convex: ears are cut at the first vertex in turn); the two-vertex face is dropped
    want = [(0, 1, 2), (0, 1, 3), (1, 2, 3), (0, 1, 2), (0, 1, 2), (0, 2, 4), (0, 4, 3)]
    assert m["tris"].shape == (len(want), 9)
    for t, (a, b, c) in zip(m["tris"], want):
        assert np.array_equal(t, np.concatenate([P[a], P[b], P[c]]))
    assert list(m["texid"]) == [0, 1, 1, 0, 0, 0, 0]                 # unknown material -> 0 like "no material"
    ex = m["triex"]
    assert np.array_equal(ex[0, :9], np.tile([0, 0, 1], 3)) and np.array_equal(ex[0, 9:15], [0, 0, 1, 0, 1, 1])
    assert np.array_equal(ex[1, :9], np.tile([0, 1, 0], 3)) and not ex[1, 9:15].any()
    assert not ex[3].any()                                           # no normals / UVs given: zeros
    mats = m["mats_raw"]
    assert len(mats) == 2
    assert np.allclose(mats[0]["ambient"], [0.1, 0.2, 0.3]) and np.allclose(mats[0]["diffuse"], [0.9, 0.1, 0.1])
    assert np.allclose(mats[0]["emissive"], [0, 0, 1]) and mats[0]["shininess"] == 32 and mats[0]["ior"] == 1.5
    assert mats[0]["dissolve"] == 0.75 and mats[0]["illum"] == 2 and mats[0]["diffuse_tex_id"] == -1
    assert mats[1]["diffuse_tex_id"] == 0 and (mats[1]["tex_width"], mats[1]["tex_height"]) == (2, 2)
    assert np.allclose(mats[1]["diffuse"], 0.6) and mats[1]["dissolve"] == 0.75 and mats[1]["shininess"] == 1.0
    assert m["texture_names"] == ["checker.png"]
    assert np.array_equal(m["materials"][1]["tex"], pack_rgb(tex))
    # the loaded mesh goes straight into the layout-B scene builder
    hs = host.RtxHostScene([m])
    karg = hs.setup(32, 24)
    assert hs.buffers()["tri"].nbytes == len(want) * 36 and int(karg["dst_width"][0]) == 32
    with pytest.raises(RuntimeError):
        host.load_obj(str(tmp_path / "missing.obj"))


# ---------------------------------------------------------------------------------------------------- vs the reference
def _ref():
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libref_assets.so not built (needs /root/reference: make -C oracle ref)")
    L = C.CDLL(REF_SO)
    L.refa_mesh_load.restype = C.c_void_p
    L.refa_mesh_load.argtypes = [C.c_char_p]
    for n in ("refa_mesh_free", "refa_png_free"):
        getattr(L, n).argtypes = [C.c_void_p]
        getattr(L, n).restype = None
    for n in ("refa_mesh_tri_count", "refa_mesh_mat_count", "refa_mesh_tex_count"):
        getattr(L, n).argtypes = [C.c_void_p]
        getattr(L, n).restype = C.c_uint32
    L.refa_sizes.argtypes = [C.c_uint32]
    L.refa_sizes.restype = C.c_uint32
    L.refa_mesh_copy.argtypes = [C.c_void_p] * 4
    L.refa_mesh_tex_dims.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.refa_mesh_tex_copy.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
    L.refa_png_load.restype = C.c_void_p
    L.refa_png_load.argtypes = [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.refa_png_copy.argtypes = [C.c_void_p, C.c_void_p]
    return L


def _ref_mesh(L, path):
    h = L.refa_mesh_load(os.fsencode(path))
    try:
        assert (L.refa_sizes(0), L.refa_sizes(1), L.refa_sizes(2)) == (36, 64, 88)
        n, nm, nt = L.refa_mesh_tri_count(h), L.refa_mesh_mat_count(h), L.refa_mesh_tex_count(h)
        tri, ex, mats = np.zeros((n, 9), np.float32), np.zeros((n, 16), np.uint32), np.zeros(max(nm, 1), host.MATERIAL_DT)
        L.refa_mesh_copy(h, tri.ctypes.data, ex.ctypes.data, mats.ctypes.data)
        texs = []
        for i in range(nt):
            w, hh = C.c_uint32(), C.c_uint32()
            L.refa_mesh_tex_dims(h, i, C.byref(w), C.byref(hh))
            t = np.zeros((hh.value, w.value), np.uint32)
            L.refa_mesh_tex_copy(h, i, t.ctypes.data)
            texs.append(t)
        return tri, ex, mats[:nm], texs
    finally:
        L.refa_mesh_free(h)


def test_shipped_objs_match_reference_loader():
    """every .obj the reference ships (triangles, quads and the 32-gon caps of cone / cylinder): identical arrays"""
    L = _ref()
    objs = sorted(glob.glob(os.path.join(REF_ASSETS, "*.obj")))
    if not objs:
        pytest.skip("no shipped assets next to oracle/_ref")
    for path in objs:
        tri, ex, mats, _ = _ref_mesh(L, path)
        m = host.load_obj(path)
        assert m["tris"].shape == tri.shape, path
        assert m["tris"].tobytes() == tri.tobytes(), path
        assert m["triex"].tobytes() == ex[:, :15].tobytes(), path
        assert np.array_equal(m["texid"], ex[:, 15]), path
        assert len(m["mats_raw"]) == len(mats), path


def test_obj_with_materials_matches_reference_loader(tmp_path):
    L = _ref()
    (tmp_path / "scene.obj").write_text(OBJ)
    (tmp_path / "scene.mtl").write_text(MTL)
    tex = np.random.default_rng(1).integers(0, 256, size=(8, 4, 3), dtype=np.uint8)
    (tmp_path / "checker.png").write_bytes(png_bytes(tex, 2))
    tri, ex, mats, texs = _ref_mesh(L, str(tmp_path / "scene.obj"))
    m = host.load_obj(str(tmp_path / "scene.obj"))
    assert m["tris"].tobytes() == tri.tobytes() and m["triex"].tobytes() == ex[:, :15].tobytes()
    assert np.array_equal(m["texid"], ex[:, 15])
    for k in ("ambient", "diffuse", "specular", "emissive", "shininess", "ior", "dissolve", "diffuse_tex_id", "tex_width", "tex_height"):
        assert np.array_equal(m["mats_raw"][k], mats[k]), k            # the fields the reference's loader sets
    assert len(texs) == 1 and np.array_equal(m["materials"][1]["tex"], texs[0])


def test_shipped_pngs_match_reference_loader():
    L = _ref()
    pngs = sorted(glob.glob(os.path.join(REF_ASSETS, "*.png")))
    if not pngs:
        pytest.skip("no shipped assets next to oracle/_ref")
    for path in pngs:
        w, h = C.c_uint32(), C.c_uint32()
        s = L.refa_png_load(os.fsencode(path), C.byref(w), C.byref(h))
        want = np.zeros((h.value, w.value), np.uint32)
        L.refa_png_copy(s, want.ctypes.data)
        L.refa_png_free(s)
        assert np.array_equal(host.load_png(path), want), path
