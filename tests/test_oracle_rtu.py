"""CPU tests of the RT-unit (layout B) oracles — no GPU.

  * the plain-C port (oracle/port/rtu_port.c) against the golden fixtures tests/golden/rtu_*.npz, which hold outputs of the
    REFERENCE's own code (sim/simx/rt_traversal.cpp compiled verbatim) on the reference builder's own buffers;
  * the port against the compiled reference itself (oracle/_ref/libref_rtu.so) when it is present: hit records bit for bit
    on every ray that does not reach the reference's 2^32-iteration `size()-1` wrap (rt_traversal.cpp:81), frames where
    such rays are few;
  * structural invariants of the reference quantizer that the CUDA decode relies on.
"""
import glob
import os

import numpy as np
import pytest

from rt_common import ROOT
from rtu_common import SCENES_B, rtu

R = rtu()
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "rtu_*.npz")))
THREADS = os.cpu_count() or 1


def load_golden(path):
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    return g, g["karg"].view(R.KARG_B_DT).copy()


def test_fixtures_exist():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_port_matches_reference_fixture(path):
    g, karg = load_golden(path)
    port = R.RtuPort()
    img = R.Image(g, karg)
    cam = int(g["camera"])
    rays = port.primary_rays(img, cam)
    hits, work, flags = port.trace_rays(img, rays, threads=THREADS, want_work=True, want_wrap_flags=True)
    assert hits.tobytes() == g["hits"].tobytes()
    assert np.array_equal(flags, g["wrap_flags"])
    frame = port.render(img, cam, threads=THREADS)
    assert np.array_equal(frame, g["image"])
    assert work["nodes"] > 0 and work["tri_tests"] > 0


@pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref/libref_rtu.so not built (needs /root/reference)")
@pytest.mark.parametrize("name", list(SCENES_B), ids=list(SCENES_B))
def test_port_equals_compiled_reference(name):
    mk, w, h, depth, cam = SCENES_B[name]
    ref, port = R.RtuRef(), R.RtuPort()
    bufs, karg = ref.build_scene(mk(), w, h, max_depth=depth)
    img = R.Image(bufs, karg)
    rays = ref.primary_rays(img, cam)
    assert rays.tobytes() == port.primary_rays(img, cam).tobytes()
    hits, flags = port.trace_rays(img, rays, want_wrap_flags=True)
    ok = flags == 0
    assert ok.sum() > 0 or len(rays) == 1
    assert ref.trace_rays(img, rays[ok], threads=THREADS).tobytes() == hits[ok].tobytes()
    if flags.sum() <= 2:     # full frames (shading + nested traceRay recursion) where the reference does not spin for long
        assert np.array_equal(ref.render(img, cam, threads=THREADS), port.render(img, cam, threads=THREADS))
    # incoherent rays through the scene box
    rng = np.random.default_rng(7)
    o = rng.uniform(-3, 3, (2000, 3)).astype(np.float32)
    if cam == 0:
        o += np.array([50, 100, 0], np.float32)
    d = rng.normal(size=(2000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rr = np.concatenate([o, d], axis=1).astype(np.float32)
    h2, f2 = port.trace_rays(img, rr, want_wrap_flags=True)
    ok2 = f2 == 0
    assert ref.trace_rays(img, rr[ok2], threads=THREADS).tobytes() == h2[ok2].tobytes()


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_port_closest_hit_is_brute_force_closest(path):
    """size-independent property: the trail/short-stack traversal returns the minimum-t triangle of a brute-force sweep"""
    g, karg = load_golden(path)
    port = R.RtuPort()
    img = R.Image(g, karg)
    hits = g["hits"].view(R.HIT_DT)
    rays = port.primary_rays(img, int(g["camera"]))
    tri = g["tri"].view(np.float32).reshape(-1, 9).astype(np.float64)
    blas = g["blas"].view(R.BLAS_B_DT)
    if len(blas) != 1:
        pytest.skip("brute force written for single-instance scenes")
    inv = blas["invTransform"][0].reshape(4, 4).astype(np.float64)
    sel = np.flatnonzero(hits["dist"] < 1e30)[::37]
    v0, e1, e2 = tri[:, 0:3], tri[:, 3:6] - tri[:, 0:3], tri[:, 6:9] - tri[:, 0:3]
    for i in sel:
        o = inv[:3, :3] @ rays[i, :3].astype(np.float64) + inv[:3, 3]
        d = inv[:3, :3] @ rays[i, 3:].astype(np.float64)
        h = np.cross(d, e2)
        a = (e1 * h).sum(1)
        with np.errstate(divide="ignore", invalid="ignore"):
            f = 1.0 / a
            s = o - v0
            u = f * (s * h).sum(1)
            q = np.cross(s, e1)
            v = f * (q @ d)
            t = f * (e2 * q).sum(1)
        okm = (np.abs(a) >= 1e-6) & (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > 1e-6)
        tmin = t[okm].min()
        assert abs(tmin - hits["dist"][i]) <= 1e-4 * max(1.0, abs(tmin)), (i, tmin, hits["dist"][i])


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_quantized_boxes_are_conservative(path):
    """origin + ldexp(q, e) of every valid child box contains that child's own origin (its true aabbMin)"""
    g, _ = load_golden(path)
    allq = g["qbvh"].view(R.QNODE_DT)[: int(g["bvh_nodes_used"])]
    offs = sorted(int(b) for b in g["blas"].view(R.BLAS_B_DT)["bvh_offset"]) + [len(allq)]
    checked = 0
    for b0, b1 in zip(offs[:-1], offs[1:]):
        q = allq[b0:b1]          # child indices are relative to the instance's first node (rt_traversal.cpp:117,92)
        for n in q[q["leafIdx"] == 0]:
            for k in range(4):
                c = n["children"][k]
                if not c["meta"] or n["leftFirst"] + k >= len(q):
                    continue
                child = q[n["leftFirst"] + k]
                e = np.array([n["ex"], n["ey"], n["ez"]], dtype=np.float64)
                lo = n["origin"].astype(np.float64) + c["q"][:3].astype(np.float64) * np.exp2(e)
                hi = n["origin"].astype(np.float64) + c["q"][3:].astype(np.float64) * np.exp2(e)
                co = child["origin"].astype(np.float64)
                assert (lo <= co + 1e-6).all() and (hi >= co - 1e-6).all()
                checked += 1
    assert checked > 0


@pytest.mark.parametrize("path", GOLDEN[:1], ids=[os.path.basename(p)[:-4] for p in GOLDEN[:1]])
def test_port_extended_modes_are_deterministic_and_consistent(path):
    """the builder-defined shadow / diffuse modes of the RT-unit layout (rtu_port.c:trace_ext_b): thread-count independent, mode 0 is
    the shipped shaders, a shadowed pixel is never brighter than the unshadowed one"""
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    karg = g["karg"].view(R.KARG_B_DT).copy()
    img = R.Image(g, karg)
    cam = int(g["camera"])
    port = R.RtuPort()
    plain = port.render(img, cam, threads=2)
    assert np.array_equal(port.render(img, cam, threads=2, mode=0), plain)
    for mode in (1, 2, 3):
        a, b = port.render(img, cam, threads=1, mode=mode), port.render(img, cam, threads=4, mode=mode)
        assert np.array_equal(a, b), mode
    lit, shadowed = port.render(img, cam, threads=4, mode=0), port.render(img, cam, threads=4, mode=1)
    for sh in (16, 8, 0):
        assert (((shadowed >> sh) & 255) <= ((lit >> sh) & 255)).all()
