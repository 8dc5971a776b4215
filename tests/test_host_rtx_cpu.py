"""CPU tests of the layout-B host builder (libvortex-b200-host.so, vxh_rtx_*): the product's replacement for the reference
raytracing host's Scene / BVH4 / quantizer / TLAS.  The reference builder is not reproducible (it reads uninitialised
cluster bounds), so the checks are semantic: the buffers are valid layout-B input for the reference's own RT-unit traversal
(oracle), which must find exactly the hits it finds on the reference-built buffers of the same meshes."""
import ctypes as C
import importlib
import os

import numpy as np
import pytest

from rt_common import scenes
from rtu_common import SCENES_B, b_mesh, rtu

host = importlib.import_module("vortex-raytracing_b200.host")
R = rtu()


def test_exports():
    L = C.CDLL(host.LIB_PATH)
    for n in host.RTX_EXPORTS:
        getattr(L, n)


@pytest.mark.parametrize("name", list(SCENES_B), ids=list(SCENES_B))
def test_buffers_are_valid_rtu_input(name):
    mk, w, h, depth, cam = SCENES_B[name]
    meshes = mk()
    hs = host.RtxHostScene(meshes)
    karg = hs.setup(w, h, max_depth=depth)
    bufs = hs.buffers()
    n_tri = sum(m["tris"].shape[0] for m in meshes)
    assert len(bufs["tri"]) == 36 * n_tri and len(bufs["triEx"]) == 64 * n_tri and len(bufs["qbvh"]) == 52 * 2 * n_tri
    # the triangle buffer is a permutation of the input, mesh by mesh, and triEx moved with it
    src = np.concatenate([m["tris"].reshape(-1, 9) for m in meshes]).astype(np.float32)
    out = bufs["tri"].view(np.float32).reshape(-1, 9)
    assert sorted(map(bytes, src)) == sorted(map(bytes, out))
    q = bufs["qbvh"].view(R.QNODE_DT)[: bufs["bvh_nodes_used"]]
    assert (q["imask"] == 0).all()
    leaves = q[q["leafIdx"] != 0]
    assert int(leaves["leafIdx"].sum()) == n_tri          # every triangle in exactly one leaf
    t = bufs["tlas"].view(R.QNODE_DT)[: bufs["tlas_nodes_used"]]
    assert (t["imask"] == 1).all()
    assert sorted(int(x) for x in t["leafIdx"][t["leafIdx"] != 0xFFFFFFFF]) == list(range(len(meshes)))
    # decoded child boxes contain the triangles of BLAS leaves below them (conservative quantization)
    tri = out.reshape(-1, 3, 3)
    blas = bufs["blas"].view(R.BLAS_B_DT)
    for b in blas:
        base = int(b["bvh_offset"])
        for n in q[base:base + 64]:
            if n["leafIdx"] != 0:
                continue
            for k in range(4):
                c = n["children"][k]
                if not c["meta"]:
                    continue
                child = q[base + n["leftFirst"] + k]
                if child["leafIdx"] == 0:
                    continue
                e = np.array([n["ex"], n["ey"], n["ez"]], dtype=np.int32)
                lo = n["origin"] + np.ldexp(c["q"][:3].astype(np.float32), e)
                hi = n["origin"] + np.ldexp(c["q"][3:].astype(np.float32), e)
                v = tri[child["leftFirst"]: child["leftFirst"] + child["leafIdx"]].reshape(-1, 3)
                assert (v >= lo).all() and (v <= hi).all()
    # the reference's traversal (C port, bit-identical to the compiled reference) on these buffers
    port = R.RtuPort()
    img = R.Image(bufs, karg)
    rays = port.primary_rays(img, cam)
    hits = port.trace_rays(img, rays, threads=os.cpu_count() or 1)
    if R.have_ref():
        rb, rk = R.RtuRef().build_scene(meshes, w, h, max_depth=depth)
        for f in ("camera_pos", "camera_forward", "camera_right", "camera_up", "viewplane"):
            assert np.array_equal(karg[f], rk[f]), f
        rimg = R.Image(rb, rk)
        hits_ref = port.trace_rays(rimg, port.primary_rays(rimg, cam), threads=os.cpu_count() or 1)
        # same geometry, same intersection arithmetic: identical distances; IDs name the same triangle
        assert np.array_equal(hits["dist"], hits_ref["dist"])
        hit = hits["dist"] < 1e30
        ta = bufs["tri"].view(np.float32).reshape(-1, 9)[hits["triIdx"][hit]]
        tb = rb["tri"].view(np.float32).reshape(-1, 9)[hits_ref["triIdx"][hit]]
        same = (ta == tb).all(axis=1)
        assert same.mean() >= 0.999   # the rest: equal-t ties on shared edges resolved in a different visiting order
        assert np.array_equal(hits["blasIdx"][hit], hits_ref["blasIdx"][hit])


def test_deterministic_and_threaded():
    m = [b_mesh(scenes.heightfield(48))]
    a = host.RtxHostScene(m, threads=1)
    a.setup(64, 64)
    b = host.RtxHostScene(m, threads=8)
    b.setup(64, 64)
    ba, bb = a.buffers(), b.buffers()
    for n in host.RTX_BUFFERS:
        assert np.array_equal(ba[n], bb[n]), n


def test_degenerate_inputs():
    one = scenes.heightfield(1)
    m = b_mesh({"tris": one["tris"][:1], "triex": one["triex"][:1], "tex": one["tex"], "reflectivity": 0.0})
    hs = host.RtxHostScene([m])
    karg = hs.setup(8, 8)
    bufs = hs.buffers()
    assert bufs["bvh_nodes_used"] == 1 and bufs["tlas_nodes_used"] == 1
    port = R.RtuPort()
    img = R.Image(bufs, karg)
    c = m["tris"][0].reshape(3, 3).mean(axis=0)
    rays = np.array([[c[0], c[1] + 2.0, c[2], 0.0, -1.0, 0.0], [c[0], c[1] + 2.0, c[2], 0.0, 1.0, 0.0]], dtype=np.float32)
    hits = port.trace_rays(img, rays)
    assert hits["dist"][0] < 1e30 and hits["triIdx"][0] == 0 and hits["dist"][1] == np.float32(1e30)
