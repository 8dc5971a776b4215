"""CPU tests of the oracle (no GPU): the C port of the reference's raycast path is pinned against
  (1) the golden fixtures of tests/golden/ — outputs of the reference's own code (see tests/golden/make_golden.py), and
  (2) the compiled reference itself (oracle/_ref/libref_raycast.so) when it is present, bit for bit.
The reference ships no golden vectors for this path; these are what pins parity."""
import glob
import hashlib
import os
import subprocess

import numpy as np
import pytest

from rt_common import ROOT, pyoracle, scenes

po = pyoracle()
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "ref_*.npz")))


@pytest.fixture(scope="module")
def port():
    if not po.have_port():
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, stdout=subprocess.DEVNULL)
    return po.PortRaycast()


def load_golden(path):
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    g["karg"] = g["karg"].view(po.KARG_DT).copy()
    g["hits"] = g["hits"].view(po.HIT_DT).copy()
    return g


def test_golden_present():
    assert len(GOLDEN) >= 4


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_port_hot_path_on_golden_buffers(port, path):
    """traversal + intersection + shading of the port, on the reference-built buffers, equals the reference's outputs"""
    g = load_golden(path)
    rs = po.RawScene(port, g, g["karg"])
    img, _ = rs.render(mode=int(g["mode"]), threads=2)
    assert np.array_equal(img, g["image"])
    hits = rs.trace_primary(threads=2)
    assert hits.tobytes() == g["hits"].tobytes()      # dist, bcoords, blasIdx, triIdx: bit-exact


@pytest.mark.parametrize("name", ["ref_cornell_64", "ref_multi_d3_96"])
def test_port_builders_reproduce_reference_buffers(port, name):
    """the port's BVH2 / TLAS builders and Tracer::setup produce the reference's upload bytes for procedural scenes"""
    g = load_golden(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    if name == "ref_cornell_64":
        meshes, (w, h, d) = [scenes.cornell_box()], (64, 64, 1)
    else:
        meshes = [scenes.heightfield(16), scenes.icosphere((0, 0.5, 0), 0.4, 2, reflectivity=0.5), scenes.cornell_box(0.3)]
        w, h, d = 96, 72, 3
    sc = port.scene_create(meshes)
    karg = sc.setup(w, h, spp=1, max_depth=d)
    b = sc.buffers()
    for n in po.BUFFER_NAMES:
        assert b[n].tobytes() == g[n].tobytes(), n
    assert b["bvh_nodes_used"] == int(g["bvh_nodes_used"])
    for f in karg.dtype.names:
        if not f.endswith("_addr"):
            assert karg[f].tobytes() == g["karg"][f].tobytes(), f


def test_port_builder_on_asset_triangles(port):
    """teapot triangles (from the golden fixture, loaded by the reference's OBJ loader) -> same BVH bytes"""
    g = load_golden(os.path.join(ROOT, "tests", "golden", "ref_teapot3_d4_64.npz"))
    tri = g["tri"].view(np.float32).reshape(-1, 9)
    triex = g["triEx"].view(np.float32).reshape(-1, 15)
    counts = g["tri_counts"]
    blas = g["blas"].view(np.uint8).reshape(-1, 160)
    meshes, off, toff = [], 0, 0
    for i, c in enumerate(counts):
        tw, th = blas[i, 144:152].view(np.uint32)
        refl = blas[i, 152:156].view(np.float32)[0]
        tex = g["tex"][toff:toff + int(tw) * int(th) * 4].view(np.uint32).reshape(int(th), int(tw))
        meshes.append({"tris": tri[off:off + c], "triex": triex[off:off + c], "tex": tex, "reflectivity": float(refl)})
        off += int(c)
        toff += int(tw) * int(th) * 4
    sc = port.scene_create(meshes)
    b = sc.buffers()
    assert b["bvh"].tobytes() == g["bvh"].tobytes()
    assert b["triIdx"].tobytes() == g["triIdx"].tobytes()
    assert b["bvh_nodes_used"] == int(g["bvh_nodes_used"])


CASES = [
    ("cornell", lambda: [scenes.cornell_box()], 48, 48, 1, 0),
    ("cornell_mirror", lambda: [scenes.cornell_box(0.5)], 64, 48, 4, 0),
    ("hf_shadow", lambda: [scenes.heightfield(48)], 96, 64, 1, 1),
    ("hf_diffuse", lambda: [scenes.heightfield(24)], 64, 64, 4, 2),
    ("multi_both", lambda: [scenes.heightfield(12), scenes.icosphere((0.2, 0.4, 0), 0.35, 1, reflectivity=0.4)], 80, 60, 3, 3),
]


@pytest.mark.skipif(not po.have_ref(), reason="compiled reference (oracle/_ref) not present")
@pytest.mark.parametrize("name,mk,w,h,depth,mode", CASES, ids=[c[0] for c in CASES])
def test_port_equals_compiled_reference(port, name, mk, w, h, depth, mode):
    ref = po.RefRaycast()
    meshes = mk()
    a, b = ref.scene_create(meshes), port.scene_create(meshes)
    ka, kb = a.setup(w, h, max_depth=depth), b.setup(w, h, max_depth=depth)
    ba, bb = a.buffers(), b.buffers()
    for n in po.BUFFER_NAMES:
        assert ba[n].tobytes() == bb[n].tobytes(), n
    ia, na = a.render(mode=mode, threads=2)
    ib, nb = b.render(mode=mode, threads=2)
    assert np.array_equal(ia, ib) and na == nb
    assert a.trace_primary().tobytes() == b.trace_primary().tobytes()
    rays = scenes.random_rays(4096, (-1.2, -0.5, -1.2), (1.2, 1.0, 1.2), seed=3)
    assert a.trace_rays(rays).tobytes() == b.trace_rays(rays).tobytes()


def test_port_edge_rays(port):
    """axis-parallel rays (zero direction components => 0*inf = NaN in the slab test), rays starting on box planes,
    rays grazing shared edges: the port must agree with itself across thread counts and never crash; when the compiled
    reference is present it must agree with it bit for bit."""
    sc = port.scene_create([scenes.cornell_box()])
    sc.setup(8, 8, rotate_scene=False)
    o = np.array([[0, 0, -3], [-1, 0, -3], [1, 1, -3], [0, -1, -3], [0.25, 0.999, -3], [0, 0, 0], [-1, -1, -1]], dtype=np.float32)
    d = np.array([[0, 0, 1], [0, 0, 1], [0, 0, 1], [0, 0, 1], [0, 0, 1], [0, 1, 0], [1, 1, 1]], dtype=np.float32)
    d[-1] /= np.float32(np.sqrt(3.0))
    rays = np.concatenate([o, d], axis=1)
    h1 = sc.trace_rays(rays, threads=1)
    h2 = sc.trace_rays(rays, threads=3)
    assert h1.tobytes() == h2.tobytes()
    if po.have_ref():
        rsc = po.RefRaycast().scene_create([scenes.cornell_box()])
        rsc.setup(8, 8, rotate_scene=False)
        assert rsc.trace_rays(rays).tobytes() == h1.tobytes()


def test_counters_consistent(port):
    sc = port.scene_create([scenes.heightfield(16)])
    sc.setup(32, 32)
    img, cnt = sc.render_counted(mode=0)
    img2, _ = sc.render(mode=0)
    assert np.array_equal(img, img2)
    assert cnt["rays"] == 32 * 32 and cnt["blas_entered"] == cnt["rays"]     # single instance: TLAS root is a leaf
    assert cnt["box_tests"] % 2 == 0 and cnt["tri_tests"] > 0 and cnt["shaded_hits"] <= cnt["rays"]


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "raycast")) or not os.path.exists("/root/reference"),
                    reason="unmodified reference binary or its assets not present")
def test_reference_binary_md5(tmp_path):
    """build self-check of the unmodified reference app"""
    out = tmp_path / "o.ppm"
    subprocess.run([os.path.join(ROOT, "oracle", "_ref", "raycast"), "-c", "-w64", "-h64", "-o", str(out)], check=True,
                   stdout=subprocess.DEVNULL)
    assert hashlib.md5(out.read_bytes()).hexdigest() == "a8248b0896d09d1b3b98b67f5c43ce6c"
