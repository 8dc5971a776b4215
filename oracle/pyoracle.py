"""oracle/pyoracle.py — TEST INFRASTRUCTURE ONLY (ctypes bindings to the CPU oracles).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product (vortex-raytracing_b200/) never does.

Two oracles are bound with the same Python surface:
  * RefRaycast  — oracle/_ref/libref_raycast.so: the REFERENCE's own raycast sources (render.h,
                  bvh.cpp, scene.cpp ...) compiled from /root/reference by oracle/Makefile.
  * PortRaycast — oracle/_build/liboracle_port.so: our plain-C restatement (oracle/port/rc_port.c).
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libref_raycast.so")
PORT_SO = os.path.join(HERE, "_build", "liboracle_port.so")

KERNEL_ARG_BYTES = 192  # raycast/common.h:120-147

# numpy views of the reference's layout-A records
HIT_DT = np.dtype([("dist", "<f4"), ("bx", "<f4"), ("by", "<f4"), ("bz", "<f4"), ("blasIdx", "<u4"), ("triIdx", "<u4")])
KARG_DT = np.dtype([
    ("dst_width", "<u4"), ("dst_height", "<u4"), ("dst_addr", "<u8"),
    ("tri_addr", "<u8"), ("triEx_addr", "<u8"), ("triIdx_addr", "<u8"), ("tex_addr", "<u8"),
    ("bvh_addr", "<u8"), ("blas_addr", "<u8"), ("tlas_addr", "<u8"), ("tlas_root", "<u4"),
    ("camera_pos", "<f4", 3), ("camera_forward", "<f4", 3), ("camera_right", "<f4", 3), ("camera_up", "<f4", 3),
    ("viewplane", "<f4", 2), ("samples_per_pixel", "<u4"), ("max_depth", "<u4"),
    ("light_pos", "<f4", 3), ("light_color", "<f4", 3), ("ambient_color", "<f4", 3), ("background_color", "<f4", 3),
    ("_pad", "<u4"),
])
assert KARG_DT.itemsize == KERNEL_ARG_BYTES, KARG_DT.itemsize

MODE_SHADOW = 1
MODE_DIFFUSE = 2

# defaults of the reference app (raycast/main.cpp:33-42)
DEFAULT_LIGHT_POS = (0.0, 10.0, -10.0)
DEFAULT_LIGHT_COLOR = (1.0, 1.0, 1.0)
DEFAULT_AMBIENT = (0.4, 0.4, 0.4)
DEFAULT_BACKGROUND = (0.4, 0.35, 0.25)


class _View(C.Structure):
    _fields_ = [(n, t) for pair in (
        [("tri", C.c_void_p), ("tri_bytes", C.c_uint64)],
        [("triEx", C.c_void_p), ("triEx_bytes", C.c_uint64)],
        [("triIdx", C.c_void_p), ("triIdx_bytes", C.c_uint64)],
        [("bvh", C.c_void_p), ("bvh_bytes", C.c_uint64)],
        [("blas", C.c_void_p), ("blas_bytes", C.c_uint64)],
        [("tlas", C.c_void_p), ("tlas_bytes", C.c_uint64)],
        [("tex", C.c_void_p), ("tex_bytes", C.c_uint64)],
    ) for (n, t) in pair] + [("tlas_root", C.c_uint32), ("bvh_nodes_used", C.c_uint32)]


BUFFER_NAMES = ("tri", "triEx", "triIdx", "bvh", "blas", "tlas", "tex")


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


class _Raycast:
    """Common surface of the two oracles (prefix = 'ref' or 'port')."""

    def __init__(self, so_path, prefix):
        if not os.path.exists(so_path):
            raise FileNotFoundError(f"{so_path} missing: run `make -C oracle` (or __graft_entry__.build())")
        self.lib = C.CDLL(so_path)
        self.p = prefix
        L = self.lib
        g = lambda n: getattr(L, f"{prefix}_{n}")
        g("scene_create").restype = C.c_void_p
        g("scene_create").argtypes = [C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p]
        g("scene_view").argtypes = [C.c_void_p, C.POINTER(_View)]
        g("scene_free").argtypes = [C.c_void_p]
        g("scene_build").argtypes = [C.c_void_p]
        g("scene_apply_transform").argtypes = [C.c_void_p, C.c_void_p]
        g("setup").argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float,
                               C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        g("render").restype = C.c_uint64
        g("render").argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
        g("trace_rays").argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
        g("trace_primary").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        g("hw_threads").restype = C.c_int
        self._g = g

    def hw_threads(self):
        return int(self._g("hw_threads")())

    # -- scenes -------------------------------------------------------------------------------
    def scene_create(self, meshes):
        """meshes: list of dict(tris (n,9) f32, triex (n,15) f32, tex (h,w) u32, reflectivity, transform (4,4)|None)."""
        n = len(meshes)
        counts = np.array([m["tris"].shape[0] for m in meshes], dtype=np.uint32)
        tris = np.ascontiguousarray(np.concatenate([m["tris"].reshape(-1, 9) for m in meshes]), dtype=np.float32)
        triex = np.ascontiguousarray(np.concatenate([m["triex"].reshape(-1, 15) for m in meshes]), dtype=np.float32)
        tw = np.array([m["tex"].shape[1] for m in meshes], dtype=np.uint32)
        th = np.array([m["tex"].shape[0] for m in meshes], dtype=np.uint32)
        texels = np.ascontiguousarray(np.concatenate([m["tex"].reshape(-1) for m in meshes]), dtype=np.uint32)
        refl = np.array([m.get("reflectivity", 0.0) for m in meshes], dtype=np.float32)
        have_t = any(m.get("transform") is not None for m in meshes)
        tr = None
        if have_t:
            tr = np.stack([np.asarray(m["transform"], dtype=np.float32) if m.get("transform") is not None
                           else np.eye(4, dtype=np.float32) for m in meshes]).astype(np.float32)
            tr = np.ascontiguousarray(tr)
        h = self._g("scene_create")(n, counts.ctypes.data, tris.ctypes.data, triex.ctypes.data, tw.ctypes.data,
                                    th.ctypes.data, texels.ctypes.data, refl.ctypes.data,
                                    tr.ctypes.data if tr is not None else None)
        if not h:
            raise RuntimeError("scene_create failed")
        return Scene(self, h)

    def scene_load_assets(self, model="teapot.obj", mesh_count=1):
        if self.p != "ref":
            raise RuntimeError("asset loading exists only in the reference harness")
        f = self.lib.ref_scene_load_assets
        f.restype = C.c_void_p
        f.argtypes = [C.c_char_p, C.c_uint32]
        h = f(model.encode(), mesh_count)
        if not h:
            raise RuntimeError("ref_scene_load_assets failed")
        return Scene(self, h)


class Scene:
    def __init__(self, oracle, handle):
        self.o = oracle
        self.h = handle
        self.karg = None

    def __del__(self):
        try:
            if self.h:
                self.o._g("scene_free")(self.h)
                self.h = None
        except Exception:
            pass

    def apply_transform(self, m):
        m = np.ascontiguousarray(np.asarray(m, dtype=np.float32).reshape(16))
        self.o._g("scene_apply_transform")(self.h, m.ctypes.data)

    def setup(self, width, height, spp=1, max_depth=1, vfov=45.0, zoom=1.0, rotate_scene=True,
              light_pos=DEFAULT_LIGHT_POS, light_color=DEFAULT_LIGHT_COLOR, ambient=DEFAULT_AMBIENT,
              background=DEFAULT_BACKGROUND):
        """Tracer::setup (raycast/tracer.cpp:152-196).  Returns the 192-byte kernel_arg_t (host pointers inside)."""
        buf = np.zeros(1, dtype=KARG_DT)
        self.o._g("setup")(self.h, buf.ctypes.data, width, height, spp, max_depth, vfov, zoom, int(rotate_scene),
                           _f3(light_pos), _f3(light_color), _f3(ambient), _f3(background))
        self.karg = buf
        return buf

    def view(self):
        v = _View()
        self.o._g("scene_view")(self.h, C.byref(v))
        return v

    def buffers(self):
        """Copies of the seven scene buffers, byte for byte what the reference host would upload."""
        v = self.view()
        out = {}
        for n in BUFFER_NAMES:
            nbytes = getattr(v, n + "_bytes")
            ptr = getattr(v, n)
            out[n] = np.frombuffer(C.string_at(ptr, nbytes), dtype=np.uint8).copy() if nbytes else np.zeros(0, np.uint8)
        out["tlas_root"] = int(v.tlas_root)
        out["bvh_nodes_used"] = int(v.bvh_nodes_used)
        return out

    # -- the hot path -------------------------------------------------------------------------
    def render(self, karg=None, mode=0, rows=None, threads=1):
        karg = self.karg if karg is None else karg
        w, h = int(karg["dst_width"][0]), int(karg["dst_height"][0])
        out = np.zeros((h, w), dtype=np.uint32)
        y0, y1 = (0, h) if rows is None else rows
        nrays = self.o._g("render")(karg.ctypes.data, out.ctypes.data, mode, y0, y1, threads)
        return out, int(nrays)

    def render_counted(self, karg=None, mode=0, threads=1):
        """render + work counters {rays, box_tests, tri_tests, blas_entered, shaded_hits} (oracle port only)."""
        if self.o.p != "port":
            raise RuntimeError("work counters exist only in the C port (the reference sources are compiled unmodified)")
        karg = self.karg if karg is None else karg
        w, h = int(karg["dst_width"][0]), int(karg["dst_height"][0])
        out = np.zeros((h, w), dtype=np.uint32)
        cnt = np.zeros(5, dtype=np.uint64)
        f = self.o.lib.port_render_counted
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]
        f.restype = None
        f(karg.ctypes.data, out.ctypes.data, mode, 0, h, threads, cnt.ctypes.data)
        return out, dict(zip(("rays", "box_tests", "tri_tests", "blas_entered", "shaded_hits"), (int(x) for x in cnt)))

    def trace_primary(self, karg=None, want_rays=False, threads=1):
        karg = self.karg if karg is None else karg
        w, h = int(karg["dst_width"][0]), int(karg["dst_height"][0])
        hits = np.zeros(w * h, dtype=HIT_DT)
        rays = np.zeros((w * h, 6), dtype=np.float32) if want_rays else None
        self.o._g("trace_primary")(karg.ctypes.data, hits.ctypes.data, rays.ctypes.data if want_rays else None, threads)
        return (hits, rays) if want_rays else hits

    def trace_rays(self, rays, karg=None, threads=1):
        karg = self.karg if karg is None else karg
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        hits = np.zeros(rays.shape[0], dtype=HIT_DT)
        self.o._g("trace_rays")(karg.ctypes.data, rays.ctypes.data, rays.shape[0], hits.ctypes.data, threads)
        return hits


class RawScene:
    """The hot path of an oracle run on caller-held buffers (the bytes a host uploaded), not on a scene it built:
    kernel_arg_t address fields are pointed at the numpy arrays, as the reference's use_cpu_ path does
    (raycast/tracer.cpp:102-109)."""

    _ADDR = {"tri": "tri_addr", "triEx": "triEx_addr", "triIdx": "triIdx_addr", "tlas": "tlas_addr", "blas": "blas_addr",
             "bvh": "bvh_addr", "tex": "tex_addr"}

    def __init__(self, oracle, buffers, karg):
        self.o = oracle
        self.bufs = {n: np.ascontiguousarray(buffers[n]).view(np.uint8).copy() for n in BUFFER_NAMES}
        self.karg = np.array(karg, dtype=KARG_DT).reshape(1).copy()
        for n, f in self._ADDR.items():
            self.karg[f] = self.bufs[n].ctypes.data
        self.karg["dst_addr"] = 0

    render = Scene.render
    trace_primary = Scene.trace_primary
    trace_rays = Scene.trace_rays

    def render_counted(self, mode=0, threads=1):
        return Scene.render_counted(self, mode=mode, threads=threads)


def RefRaycast():
    return _Raycast(REF_SO, "ref")


def PortRaycast():
    return _Raycast(PORT_SO, "port")


def have_ref():
    return os.path.exists(REF_SO)


def have_port():
    return os.path.exists(PORT_SO)


def best():
    """The strongest checker available: the compiled reference when present, else the C port."""
    return RefRaycast() if have_ref() else PortRaycast()
