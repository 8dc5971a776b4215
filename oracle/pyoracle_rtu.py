"""oracle/pyoracle_rtu.py — TEST INFRASTRUCTURE ONLY (ctypes bindings to the RT-unit / layout-B CPU oracles).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.

  * RtuRef  — oracle/_ref/libref_rtu.so: sim/simx/rt_traversal.cpp compiled verbatim behind a dcache_read shim, the
              reference raytracing host's Scene/BVH4/quantizer/TLAS, and the kernel.cpp / closest.cpp / miss.cpp program
              driven with the reference headers' own arithmetic (oracle/ref_harness/ref_rtu*.cpp).
  * RtuPort — oracle/_build/liboracle_port.so: the plain-C restatement (oracle/port/rtu_port.c), with work counters.

Both run on a flat memory image (`Image`): the bytes a host uploaded, at the Vortex addresses it was given.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libref_rtu.so")
PORT_SO = os.path.join(HERE, "_build", "liboracle_port.so")

HIT_DT = np.dtype([("dist", "<f4"), ("bx", "<f4"), ("by", "<f4"), ("bz", "<f4"), ("blasIdx", "<u4"), ("triIdx", "<u4")])
KARG_B_DT = np.dtype([   # raytracing/common.h:164-195
    ("dst_width", "<u4"), ("dst_height", "<u4"), ("dst_addr", "<u8"),
    ("tri_addr", "<u8"), ("triEx_addr", "<u8"), ("triIdx_addr", "<u8"), ("mat_addr", "<u8"), ("tex_addr", "<u8"),
    ("bvh_addr", "<u8"), ("qBvh_addr", "<u8"), ("blas_addr", "<u8"), ("tlas_addr", "<u8"), ("tlas_root", "<u4"),
    ("camera_pos", "<f4", 3), ("camera_forward", "<f4", 3), ("camera_right", "<f4", 3), ("camera_up", "<f4", 3),
    ("viewplane", "<f4", 2), ("samples_per_pixel", "<u4"), ("max_depth", "<u4"),
    ("light_pos", "<f4", 3), ("light_color", "<f4", 3), ("ambient_color", "<f4", 3), ("background_color", "<f4", 3),
    ("_pad", "<u4"), ("sbt_addr", "<u8"),
])
assert KARG_B_DT.itemsize == 216
MATERIAL_DT = np.dtype([  # raytracing/common.h:20-36
    ("ambient", "<f4", 3), ("diffuse", "<f4", 3), ("specular", "<f4", 3), ("emissive", "<f4", 3),
    ("shininess", "<f4"), ("ior", "<f4"), ("dissolve", "<f4"), ("reflectivity", "<f4"),
    ("diffuse_tex_id", "<i4"), ("illum", "<i4"), ("tex_width", "<u4"), ("tex_height", "<u4"), ("tex_offset", "<u8"),
])
assert MATERIAL_DT.itemsize == 88
QNODE_DT = np.dtype([   # raytracing/common.h:58-67 == sim/simx/rt_traversal.h:19-33
    ("origin", "<f4", 3), ("ex", "i1"), ("ey", "i1"), ("ez", "i1"), ("imask", "u1"), ("leftFirst", "<u4"), ("leafIdx", "<u4"),
    ("children", [("meta", "u1"), ("q", "u1", 6)], 4),
])
assert QNODE_DT.itemsize == 52
BLAS_B_DT = np.dtype([  # raytracing/common.h:86-99
    ("bvh_offset", "<u4"), ("invTransform", "<f4", 16), ("transform", "<f4", 16), ("_p0", "<u4"), ("mat_offset", "<u8"),
    ("tex_width", "<u4"), ("tex_height", "<u4"), ("reflectivity", "<f4"), ("_p1", "<u4"),
])
assert BLAS_B_DT.itemsize == 160

BUFFER_NAMES = ("tri", "triEx", "triIdx", "mat", "tex", "bvh", "qbvh", "blas", "tlas")
_ADDR_FIELD = {"tri": "tri_addr", "triEx": "triEx_addr", "triIdx": "triIdx_addr", "mat": "mat_addr", "tex": "tex_addr",
               "bvh": "bvh_addr", "qbvh": "qBvh_addr", "blas": "blas_addr", "tlas": "tlas_addr"}

DEFAULT_LIGHT_POS = (0.0, 10.0, -10.0)      # raytracing/main.cpp:34-38
DEFAULT_LIGHT_COLOR = (1.0, 1.0, 1.0)
DEFAULT_AMBIENT = (0.4, 0.4, 0.4)
DEFAULT_BACKGROUND = (0.4, 0.35, 0.25)


class Image:
    """Flat device-memory image: buffers laid out from 0x10000 up, 256-byte aligned, like the backend's arena would."""

    def __init__(self, buffers, karg):
        self.base = 0x10000
        addr = self.base
        self.addr = {}
        chunks = []
        for n in BUFFER_NAMES:
            b = np.ascontiguousarray(buffers[n]).view(np.uint8).reshape(-1)
            self.addr[n] = addr
            pad = (-len(b)) % 256 + (256 if len(b) == 0 else 0)
            chunks += [b, np.zeros(pad, np.uint8)]
            addr += len(b) + pad
        self.mem = np.concatenate(chunks) if chunks else np.zeros(0, np.uint8)
        self.karg = np.array(karg, dtype=KARG_B_DT).reshape(1).copy()
        for n, f in _ADDR_FIELD.items():
            self.karg[f] = self.addr[n]
        self.karg["dst_addr"] = 0
        self.karg["sbt_addr"] = 0
        # the four RT-unit DCRs (raytracing/tracer.cpp:252-256): tlas, blas, qBvh, tri — truncated to 32 bits
        self.dcr = np.array([self.addr["tlas"], self.addr["blas"], self.addr["qbvh"], self.addr["tri"]], dtype=np.uint32)


class _View(C.Structure):
    _fields_ = [(n, t) for pair in ([(b, C.c_void_p), (b + "_bytes", C.c_uint64)] for b in BUFFER_NAMES) for (n, t) in pair] + \
               [("tlas_root", C.c_uint32), ("bvh_nodes_used", C.c_uint32), ("tlas_nodes_used", C.c_uint32), ("_pad", C.c_uint32)]


def pack_meshes(meshes):
    """meshes: list of dict(tris (n,9), triex (n,15), texid (n,) u32 [local material index], materials: list of
    dict(diffuse=(r,g,b), tex=(h,w) u32 | None), transform (4,4) | None).  Returns the flat arrays rtuh_scene_create takes."""
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
    tex_chunks, tex_off, k = [], 0, 0
    for m in meshes:
        for mat in m["materials"]:
            mats[k]["diffuse"] = mat.get("diffuse", (0.8, 0.8, 0.8))
            mats[k]["ambient"] = (0.1, 0.1, 0.1)
            mats[k]["dissolve"] = 1.0
            t = mat.get("tex")
            if t is not None:
                t = np.ascontiguousarray(t, dtype=np.uint32)
                mats[k]["diffuse_tex_id"] = 0
                mats[k]["tex_width"], mats[k]["tex_height"] = t.shape[1], t.shape[0]
                mats[k]["tex_offset"] = tex_off
                tex_chunks.append(t.reshape(-1).view(np.uint8))
                tex_off += t.size * 4
            else:
                mats[k]["diffuse_tex_id"] = -1
            k += 1
    tex = np.concatenate(tex_chunks) if tex_chunks else np.zeros(0, np.uint8)
    tr = None
    if any(m.get("transform") is not None for m in meshes):
        tr = np.ascontiguousarray(np.stack([np.asarray(m["transform"], np.float32) if m.get("transform") is not None
                                            else np.eye(4, dtype=np.float32) for m in meshes]), dtype=np.float32)
    return counts, tris, triex, n_mats, mats, np.ascontiguousarray(tex), tr


def from_layout_a(mesh, diffuse=None):
    """A layout-A mesh dict (vortex-raytracing_b200/scenes.py) as a layout-B mesh: one material carrying its texture."""
    mat = {"tex": mesh["tex"]} if diffuse is None else {"diffuse": diffuse}
    return {"tris": mesh["tris"], "triex": mesh["triex"], "texid": np.zeros(mesh["tris"].shape[0], np.uint32),
            "materials": [mat], "transform": mesh.get("transform"), "reflectivity": mesh.get("reflectivity", 0.0)}


class RtuRef:
    """The compiled reference (libref_rtu.so)."""
    name = "ref"

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(f"{REF_SO} missing: run `make -C oracle ref` where /root/reference exists")
        L = self.lib = C.CDLL(REF_SO)
        L.rtuh_scene_create.restype = C.c_void_p
        L.rtuh_scene_create.argtypes = [C.c_uint32] + [C.c_void_p] * 6 + [C.c_uint64, C.c_void_p]
        L.rtuh_scene_build.argtypes = [C.c_void_p]
        L.rtuh_scene_view.argtypes = [C.c_void_p, C.POINTER(_View)]
        L.rtuh_scene_destroy.argtypes = [C.c_void_p]
        L.rtuh_scene_set_reflectivity.argtypes = [C.c_void_p, C.c_uint32, C.c_float]
        L.rtuh_camera.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_void_p]
        L.rtu_trace_rays.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                     C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p]
        L.rtuh_render.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
        L.rtuh_primary_rays.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]

    # ---- scene building with the reference's own (non-deterministic) BVH4 builder + quantizer ----
    def build_scene(self, meshes, width, height, spp=1, max_depth=1, vfov=45.0, zoom=1.0, light_pos=DEFAULT_LIGHT_POS,
                    light_color=DEFAULT_LIGHT_COLOR, ambient=DEFAULT_AMBIENT, background=DEFAULT_BACKGROUND):
        """Returns (buffers dict of uint8 arrays, karg record) exactly as raytracing/tracer.cpp would upload them."""
        counts, tris, triex, n_mats, mats, tex, tr = pack_meshes(meshes)
        L = self.lib
        h = L.rtuh_scene_create(len(meshes), counts.ctypes.data, tris.ctypes.data, triex.ctypes.data, n_mats.ctypes.data,
                                mats.ctypes.data if len(mats) else None, tex.ctypes.data if len(tex) else None, len(tex),
                                tr.ctypes.data if tr is not None else None)
        if not h:
            raise RuntimeError("rtuh_scene_create failed")
        for i, m in enumerate(meshes):
            if m.get("reflectivity", 0.0):
                L.rtuh_scene_set_reflectivity(h, i, float(m["reflectivity"]))
        L.rtuh_scene_build(h)
        v = _View()
        L.rtuh_scene_view(h, C.byref(v))
        bufs = {}
        for n in BUFFER_NAMES:
            nb = getattr(v, n + "_bytes")
            bufs[n] = np.frombuffer(C.string_at(getattr(v, n), nb), dtype=np.uint8).copy() if nb else np.zeros(0, np.uint8)
        bufs["bvh_nodes_used"], bufs["tlas_nodes_used"] = int(v.bvh_nodes_used), int(v.tlas_nodes_used)
        karg = np.zeros(1, dtype=KARG_B_DT)
        L.rtuh_camera(h, width, height, vfov, zoom, karg.ctypes.data)
        karg["samples_per_pixel"], karg["max_depth"] = spp, max_depth
        karg["light_pos"], karg["light_color"] = light_pos, light_color
        karg["ambient_color"], karg["background_color"] = ambient, background
        L.rtuh_scene_destroy(h)
        return bufs, karg

    def trace_rays(self, img, rays, threads=1):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        hits = np.zeros(rays.shape[0], dtype=HIT_DT)
        d = img.dcr
        self.lib.rtu_trace_rays(img.mem.ctypes.data, img.base, len(img.mem), int(d[0]), int(d[1]), int(d[2]), int(d[3]),
                                rays.ctypes.data, rays.shape[0], hits.ctypes.data, threads, None)
        return hits

    def primary_rays(self, img, camera):
        w, h = int(img.karg["dst_width"][0]), int(img.karg["dst_height"][0])
        rays = np.zeros((w * h, 6), dtype=np.float32)
        self.lib.rtuh_primary_rays(img.karg.ctypes.data, camera, rays.ctypes.data)
        return rays

    def render(self, img, camera=1, threads=1):
        w, h = int(img.karg["dst_width"][0]), int(img.karg["dst_height"][0])
        out = np.zeros((h, w), dtype=np.uint32)
        self.lib.rtuh_render(img.mem.ctypes.data, img.base, len(img.mem), img.karg.ctypes.data, img.dcr.ctypes.data, camera,
                             out.ctypes.data, threads)
        return out


WORK_NAMES = ("traverse_calls", "nodes", "blas_entered", "tri_tests", "restarts", "shaded_hits")


class RtuPort:
    """The plain-C restatement (liboracle_port.so), with work counters."""
    name = "port"

    def __init__(self):
        if not os.path.exists(PORT_SO):
            raise FileNotFoundError(f"{PORT_SO} missing: run `make -C oracle port`")
        L = self.lib = C.CDLL(PORT_SO)
        L.port_rtu_trace_rays.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                          C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.port_rtuh_render.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                       C.c_uint32, C.c_void_p]
        L.port_rtuh_primary_rays.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.port_rtu_set_fix_tlas_base.argtypes = [C.c_int]

    def set_fix_tlas_base(self, on):
        """0 (default): reproduce the reference's stale-base_ptr defect on popped TLAS inner nodes (rt_traversal.cpp:179-213);
        1: address TLAS children from the TLAS base.  See oracle/port/rtu_port.c."""
        self.lib.port_rtu_set_fix_tlas_base(int(bool(on)))

    def trace_rays(self, img, rays, threads=1, want_work=False, want_wrap_flags=False):
        """want_wrap_flags: also return one byte per ray, 1 where traversal reaches the reference's `size()-1` wrap
        (rt_traversal.cpp:81 with k == 4 and no surviving child) — the compiled reference spins 2^32 iterations there."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        hits = np.zeros(rays.shape[0], dtype=HIT_DT)
        work = np.zeros(6, dtype=np.uint64)
        flags = np.zeros(rays.shape[0], dtype=np.uint8) if want_wrap_flags else None
        d = img.dcr
        self.lib.port_rtu_trace_rays(img.mem.ctypes.data, img.base, len(img.mem), int(d[0]), int(d[1]), int(d[2]), int(d[3]),
                                     rays.ctypes.data, rays.shape[0], hits.ctypes.data, threads,
                                     work.ctypes.data if want_work else None, flags.ctypes.data if want_wrap_flags else None)
        out = (hits,)
        if want_work:
            w = dict(zip(WORK_NAMES, (int(x) for x in work)))
            w["wraps"] = w.pop("shaded_hits")
            out += (w,)
        if want_wrap_flags:
            out += (flags,)
        return out if len(out) > 1 else hits

    def primary_rays(self, img, camera):
        w, h = int(img.karg["dst_width"][0]), int(img.karg["dst_height"][0])
        rays = np.zeros((w * h, 6), dtype=np.float32)
        self.lib.port_rtuh_primary_rays(img.karg.ctypes.data, camera, rays.ctypes.data)
        return rays

    def render(self, img, camera=1, threads=1, want_work=False, mode=0):
        """mode: 0 = the shipped shaders; bit0 shadow rays, bit1 diffuse bounces (the builder-defined extension, rtu_port.c)"""
        w, h = int(img.karg["dst_width"][0]), int(img.karg["dst_height"][0])
        out = np.zeros((h, w), dtype=np.uint32)
        work = np.zeros(6, dtype=np.uint64)
        if mode:
            self.lib.port_rtuh_render_ext.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32,
                                                      C.c_void_p, C.c_uint32, C.c_void_p]
            self.lib.port_rtuh_render_ext(img.mem.ctypes.data, img.base, len(img.mem), img.karg.ctypes.data, img.dcr.ctypes.data, camera,
                                          int(mode), out.ctypes.data, threads, work.ctypes.data if want_work else None)
        else:
            self.lib.port_rtuh_render(img.mem.ctypes.data, img.base, len(img.mem), img.karg.ctypes.data, img.dcr.ctypes.data, camera,
                                      out.ctypes.data, threads, work.ctypes.data if want_work else None)
        return (out, dict(zip(WORK_NAMES, (int(x) for x in work)))) if want_work else out


def have_ref():
    return os.path.exists(REF_SO)


def best():
    return RtuRef() if have_ref() else RtuPort()
