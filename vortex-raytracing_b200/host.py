"""ctypes binding of libvortex-b200-host.so (include/vortex_b200_host.h): host-side scene preparation in the reference's
layout-A byte formats — what Scene::init / BVH / TLAS / Tracer::setup do in a Vortex raycast host
(tests/regression/raycast/scene.cpp:17-82, bvh.cpp, tracer.cpp:152-196)."""
import ctypes as C
import os

import numpy as np

from .tracer import KARG_DT, SCENE_BUFFERS

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libvortex-b200-host.so")
EXPORTS = "vxh_scene_create vxh_scene_apply_transform vxh_scene_build vxh_scene_view vxh_scene_free vxh_setup".split()

# defaults of the reference app (raycast/main.cpp:33-42)
DEFAULT_LIGHT_POS = (0.0, 10.0, -10.0)
DEFAULT_LIGHT_COLOR = (1.0, 1.0, 1.0)
DEFAULT_AMBIENT = (0.4, 0.4, 0.4)
DEFAULT_BACKGROUND = (0.4, 0.35, 0.25)


class _View(C.Structure):
    _fields_ = [(n, t) for name in ("tri", "triEx", "triIdx", "bvh", "blas", "tlas", "tex")
                for (n, t) in ((name, C.c_void_p), (name + "_bytes", C.c_uint64))] + [("tlas_root", C.c_uint32), ("bvh_nodes_used", C.c_uint32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C vortex-raytracing_b200/csrc`")
        L = C.CDLL(LIB_PATH)
        L.vxh_scene_create.restype = C.c_void_p
        L.vxh_scene_create.argtypes = [C.c_uint32] + [C.c_void_p] * 8 + [C.c_int]
        L.vxh_scene_apply_transform.argtypes = [C.c_void_p, C.c_void_p]
        L.vxh_scene_build.argtypes = [C.c_void_p]
        L.vxh_scene_view.argtypes = [C.c_void_p, C.POINTER(_View)]
        L.vxh_scene_free.argtypes = [C.c_void_p]
        L.vxh_setup.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_int,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        for n in ("vxh_scene_apply_transform", "vxh_scene_build", "vxh_scene_view", "vxh_scene_free", "vxh_setup"):
            getattr(L, n).restype = None
        _lib = L
    return _lib


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


class HostScene:
    """meshes: list of dict(tris (n,9) f32, triex (n,15) f32, tex (h,w) u32, reflectivity, transform (4,4)|None)."""

    def __init__(self, meshes, threads=0):
        n = len(meshes)
        counts = np.array([m["tris"].shape[0] for m in meshes], dtype=np.uint32)
        tris = np.ascontiguousarray(np.concatenate([m["tris"].reshape(-1, 9) for m in meshes]), dtype=np.float32)
        triex = np.ascontiguousarray(np.concatenate([m["triex"].reshape(-1, 15) for m in meshes]), dtype=np.float32)
        tw = np.array([m["tex"].shape[1] for m in meshes], dtype=np.uint32)
        th = np.array([m["tex"].shape[0] for m in meshes], dtype=np.uint32)
        texels = np.ascontiguousarray(np.concatenate([m["tex"].reshape(-1) for m in meshes]), dtype=np.uint32)
        refl = np.array([m.get("reflectivity", 0.0) for m in meshes], dtype=np.float32)
        tr = None
        if any(m.get("transform") is not None for m in meshes):
            tr = np.ascontiguousarray(np.stack([np.asarray(m["transform"], dtype=np.float32) if m.get("transform") is not None
                                                else np.eye(4, dtype=np.float32) for m in meshes]), dtype=np.float32)
        self.h = lib().vxh_scene_create(n, counts.ctypes.data, tris.ctypes.data, triex.ctypes.data, tw.ctypes.data, th.ctypes.data,
                                        texels.ctypes.data, refl.ctypes.data, tr.ctypes.data if tr is not None else None, threads)
        if not self.h:
            raise RuntimeError("vxh_scene_create failed")
        self.karg = None

    def __del__(self):
        try:
            if self.h:
                lib().vxh_scene_free(self.h)
                self.h = None
        except Exception:
            pass

    def apply_transform(self, m):
        m = np.ascontiguousarray(np.asarray(m, dtype=np.float32).reshape(16))
        lib().vxh_scene_apply_transform(self.h, m.ctypes.data)

    def setup(self, width, height, spp=1, max_depth=1, vfov=45.0, zoom=1.0, rotate_scene=True, light_pos=DEFAULT_LIGHT_POS,
              light_color=DEFAULT_LIGHT_COLOR, ambient=DEFAULT_AMBIENT, background=DEFAULT_BACKGROUND):
        buf = np.zeros(1, dtype=KARG_DT)
        lib().vxh_setup(self.h, buf.ctypes.data, width, height, spp, max_depth, vfov, zoom, int(rotate_scene), _f3(light_pos),
                        _f3(light_color), _f3(ambient), _f3(background))
        self.karg = buf
        return buf

    def buffers(self):
        v = _View()
        lib().vxh_scene_view(self.h, C.byref(v))
        out = {}
        for n in SCENE_BUFFERS:
            nbytes = getattr(v, n + "_bytes")
            out[n] = np.frombuffer(C.string_at(getattr(v, n), nbytes), dtype=np.uint8).copy() if nbytes else np.zeros(0, np.uint8)
        out["tlas_root"] = int(v.tlas_root)
        out["bvh_nodes_used"] = int(v.bvh_nodes_used)
        return out


# ---------------------------------------------------------------------------------------------------------------------
# layout B (RT-unit / "raytracing" host path): vxh_rtx_* of include/vortex_b200_host.h
RTX_EXPORTS = ("vxh_rtx_scene_create vxh_rtx_scene_apply_transform vxh_rtx_scene_build vxh_rtx_scene_view vxh_rtx_scene_free "
               "vxh_rtx_setup").split()
RTX_BUFFERS = ("tri", "triEx", "triIdx", "mat", "tex", "bvh", "qbvh", "blas", "tlas")
MATERIAL_DT = np.dtype([  # tests/regression/raytracing/common.h:20-36
    ("ambient", "<f4", 3), ("diffuse", "<f4", 3), ("specular", "<f4", 3), ("emissive", "<f4", 3),
    ("shininess", "<f4"), ("ior", "<f4"), ("dissolve", "<f4"), ("reflectivity", "<f4"),
    ("diffuse_tex_id", "<i4"), ("illum", "<i4"), ("tex_width", "<u4"), ("tex_height", "<u4"), ("tex_offset", "<u8"),
])
assert MATERIAL_DT.itemsize == 88


class _RtxView(C.Structure):
    _fields_ = [(n, t) for name in RTX_BUFFERS for (n, t) in ((name, C.c_void_p), (name + "_bytes", C.c_uint64))] + \
               [("tlas_root", C.c_uint32), ("bvh_nodes_used", C.c_uint32), ("tlas_nodes_used", C.c_uint32), ("_pad", C.c_uint32)]


def _rtx_lib():
    L = lib()
    if not getattr(L, "_rtx_ready", False):
        L.vxh_rtx_scene_create.restype = C.c_void_p
        L.vxh_rtx_scene_create.argtypes = [C.c_uint32] + [C.c_void_p] * 6 + [C.c_uint64, C.c_void_p, C.c_void_p, C.c_int]
        L.vxh_rtx_scene_apply_transform.argtypes = [C.c_void_p, C.c_void_p]
        L.vxh_rtx_scene_build.argtypes = [C.c_void_p]
        L.vxh_rtx_scene_view.argtypes = [C.c_void_p, C.POINTER(_RtxView)]
        L.vxh_rtx_scene_free.argtypes = [C.c_void_p]
        L.vxh_rtx_setup.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_float,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        for n in RTX_EXPORTS[1:]:
            getattr(L, n).restype = None
        L._rtx_ready = True
    return L


class RtxHostScene:
    """meshes: list of dict(tris (n,9) f32, triex (n,15) f32, texid (n,) u32 [material index local to the mesh],
    materials: list of dict(diffuse=(r,g,b) | tex=(h,w) u32), reflectivity, transform (4,4) | None)."""

    def __init__(self, meshes, threads=0):
        from .rtx_tracer import KARG_B_DT
        self._karg_dt = KARG_B_DT
        counts = np.array([m["tris"].shape[0] for m in meshes], dtype=np.uint32)
        tris = np.ascontiguousarray(np.concatenate([m["tris"].reshape(-1, 9) for m in meshes]), dtype=np.float32)
        ex = []
        for m in meshes:
            n = m["tris"].shape[0]
            rec = np.zeros((n, 16), dtype=np.uint32)
            rec[:, :15] = np.ascontiguousarray(m["triex"].reshape(n, 15), dtype=np.float32).view(np.uint32)
            rec[:, 15] = np.asarray(m.get("texid", np.zeros(n, np.uint32)), dtype=np.uint32)
            ex.append(rec)
        triex = np.ascontiguousarray(np.concatenate(ex))
        n_mats = np.array([len(m["materials"]) for m in meshes], dtype=np.uint32)
        mats = np.zeros(int(n_mats.sum()), dtype=MATERIAL_DT)
        chunks, off, k = [], 0, 0
        for m in meshes:
            for mat in m["materials"]:
                mats[k]["diffuse"] = mat.get("diffuse", (0.8, 0.8, 0.8))
                mats[k]["ambient"], mats[k]["dissolve"] = (0.1, 0.1, 0.1), 1.0
                t = mat.get("tex")
                if t is not None:
                    t = np.ascontiguousarray(t, dtype=np.uint32)
                    mats[k]["diffuse_tex_id"] = 0
                    mats[k]["tex_width"], mats[k]["tex_height"], mats[k]["tex_offset"] = t.shape[1], t.shape[0], off
                    chunks.append(t.reshape(-1).view(np.uint8))
                    off += t.size * 4
                else:
                    mats[k]["diffuse_tex_id"] = -1
                k += 1
        tex = np.ascontiguousarray(np.concatenate(chunks)) if chunks else np.zeros(0, np.uint8)
        refl = np.array([m.get("reflectivity", 0.0) or 0.0 for m in meshes], dtype=np.float32)
        tr = None
        if any(m.get("transform") is not None for m in meshes):
            tr = np.ascontiguousarray(np.stack([np.asarray(m["transform"], dtype=np.float32) if m.get("transform") is not None
                                                else np.eye(4, dtype=np.float32) for m in meshes]), dtype=np.float32)
        self.h = _rtx_lib().vxh_rtx_scene_create(len(meshes), counts.ctypes.data, tris.ctypes.data, triex.ctypes.data, n_mats.ctypes.data,
                                                 mats.ctypes.data if len(mats) else None, tex.ctypes.data if len(tex) else None, len(tex),
                                                 refl.ctypes.data, tr.ctypes.data if tr is not None else None, threads)
        if not self.h:
            raise RuntimeError("vxh_rtx_scene_create failed")
        self.karg = None

    def __del__(self):
        try:
            if self.h:
                _rtx_lib().vxh_rtx_scene_free(self.h)
                self.h = None
        except Exception:
            pass

    def apply_transform(self, m):
        """Scene::applyTransform (raytracing/scene.cpp:247-251)"""
        m = np.ascontiguousarray(np.asarray(m, dtype=np.float32).reshape(16))
        _rtx_lib().vxh_rtx_scene_apply_transform(self.h, m.ctypes.data)

    def setup(self, width, height, spp=1, max_depth=1, vfov=45.0, zoom=1.0, light_pos=DEFAULT_LIGHT_POS,
              light_color=DEFAULT_LIGHT_COLOR, ambient=DEFAULT_AMBIENT, background=DEFAULT_BACKGROUND):
        buf = np.zeros(1, dtype=self._karg_dt)
        _rtx_lib().vxh_rtx_setup(self.h, buf.ctypes.data, width, height, spp, max_depth, vfov, zoom, _f3(light_pos), _f3(light_color),
                                 _f3(ambient), _f3(background))
        self.karg = buf
        return buf

    def buffers(self):
        v = _RtxView()
        _rtx_lib().vxh_rtx_scene_view(self.h, C.byref(v))
        out = {}
        for n in RTX_BUFFERS:
            nbytes = getattr(v, n + "_bytes")
            out[n] = np.frombuffer(C.string_at(getattr(v, n), nbytes), dtype=np.uint8).copy() if nbytes else np.zeros(0, np.uint8)
        out["bvh_nodes_used"], out["tlas_nodes_used"] = int(v.bvh_nodes_used), int(v.tlas_nodes_used)
        return out


# ----------------------------------------------------------------------------------------------------------------------
# asset path: OBJ/MTL/PNG in, PPM out — vortex-raytracing_b200/host/asset_io.cpp
# ----------------------------------------------------------------------------------------------------------------------
ASSET_EXPORTS = ("vxh_obj_load vxh_asset_mesh_view vxh_asset_mesh_texture_name vxh_asset_mesh_free vxh_png_load vxh_free "
                 "vxh_ppm_write vxh_asset_last_error").split()


class _AssetView(C.Structure):
    _fields_ = [("tris", C.c_void_p), ("triex16", C.c_void_p), ("tri_count", C.c_uint32), ("n_mats", C.c_uint32),
                ("mats", C.c_void_p), ("tex", C.c_void_p), ("tex_bytes", C.c_uint64), ("n_textures", C.c_uint32), ("_pad", C.c_uint32)]


def _asset_lib():
    L = lib()
    if not getattr(L, "_asset_ready", False):
        L.vxh_obj_load.restype = C.c_void_p
        L.vxh_obj_load.argtypes = [C.c_char_p]
        L.vxh_asset_mesh_view.argtypes = [C.c_void_p, C.POINTER(_AssetView)]
        L.vxh_asset_mesh_view.restype = None
        L.vxh_asset_mesh_texture_name.restype = C.c_char_p
        L.vxh_asset_mesh_texture_name.argtypes = [C.c_void_p, C.c_uint32]
        L.vxh_asset_mesh_free.argtypes = [C.c_void_p]
        L.vxh_asset_mesh_free.restype = None
        L.vxh_png_load.argtypes = [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_void_p)]
        L.vxh_free.argtypes = [C.c_void_p]
        L.vxh_free.restype = None
        L.vxh_ppm_write.argtypes = [C.c_char_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int]
        L.vxh_asset_last_error.restype = C.c_char_p
        L._asset_ready = True
    return L


def _asset_error(L, what):
    raise RuntimeError(f"{what}: {L.vxh_asset_last_error().decode(errors='replace')}")


def load_obj(path, reflectivity=0.0, transform=None):
    """Mesh::Mesh(objFile) (raytracing/mesh.cpp:215-293): one mesh dict in the form RtxHostScene takes, plus the raw
    material_info_t records and texture blob under 'mats_raw' / 'tex_raw' / 'texture_names'."""
    L = _asset_lib()
    h = L.vxh_obj_load(os.fsencode(path))
    if not h:
        _asset_error(L, f"vxh_obj_load({path})")
    try:
        v = _AssetView()
        L.vxh_asset_mesh_view(h, C.byref(v))
        n = int(v.tri_count)
        tris = np.frombuffer(C.string_at(v.tris, n * 36), dtype=np.float32).reshape(n, 9).copy() if n else np.zeros((0, 9), np.float32)
        ex = np.frombuffer(C.string_at(v.triex16, n * 64), dtype=np.uint32).reshape(n, 16).copy() if n else np.zeros((0, 16), np.uint32)
        mats = np.frombuffer(C.string_at(v.mats, int(v.n_mats) * 88), dtype=MATERIAL_DT).copy() if v.n_mats else np.zeros(0, MATERIAL_DT)
        tex = np.frombuffer(C.string_at(v.tex, int(v.tex_bytes)), dtype=np.uint8).copy() if v.tex_bytes else np.zeros(0, np.uint8)
        names = [L.vxh_asset_mesh_texture_name(h, i).decode() for i in range(int(v.n_textures))]
    finally:
        L.vxh_asset_mesh_free(h)
    materials = []
    for m in mats:
        if m["diffuse_tex_id"] >= 0:
            w, hgt, off = int(m["tex_width"]), int(m["tex_height"]), int(m["tex_offset"])
            materials.append({"tex": tex[off:off + 4 * w * hgt].view(np.uint32).reshape(hgt, w), "diffuse": tuple(m["diffuse"])})
        else:
            materials.append({"diffuse": tuple(float(x) for x in m["diffuse"])})
    if not materials:   # an OBJ without a material library: every triangle has texId 0
        materials = [{"diffuse": (0.6, 0.6, 0.6)}]
    return {"tris": tris, "triex": ex[:, :15].view(np.float32).copy(), "texid": ex[:, 15].copy(), "materials": materials,
            "reflectivity": float(reflectivity), "transform": transform, "mats_raw": mats, "tex_raw": tex, "texture_names": names}


def load_png(path):
    """Surface(file) for PNG (raytracing/surface.cpp:40-52): (h, w) uint32 texels 0x00RRGGBB"""
    L = _asset_lib()
    w, h, p = C.c_uint32(), C.c_uint32(), C.c_void_p()
    if L.vxh_png_load(os.fsencode(path), C.byref(w), C.byref(h), C.byref(p)) != 0:
        _asset_error(L, f"vxh_png_load({path})")
    try:
        return np.frombuffer(C.string_at(p, w.value * h.value * 4), dtype=np.uint32).reshape(h.value, w.value).copy()
    finally:
        L.vxh_free(p)


def write_ppm(path, frame, binary=True, flip_vertical=True):
    """frame: (h, w) uint32 0x00RRGGBB as the kernels write it.  binary=False reproduces the reference's P3 text file
    (raytracing/tracer.cpp:15-33) byte for byte; binary=True is the same picture as P6."""
    L = _asset_lib()
    f = np.ascontiguousarray(frame, dtype=np.uint32)
    if L.vxh_ppm_write(os.fsencode(path), f.ctypes.data, f.shape[1], f.shape[0], 6 if binary else 3, int(bool(flip_vertical))) != 0:
        _asset_error(L, f"vxh_ppm_write({path})")
