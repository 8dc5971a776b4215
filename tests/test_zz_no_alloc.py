"""vx_b200_build_bvh_static — the GPU BVH build with its scratch in a static __device__ array (no allocation; no counterpart in the
reference, added for callers that cannot call cudaMalloc).  Through the C ABI, against vx_b200_build_bvh on the same triangles:
the bytes left in the bvh / triIdx buffers must be identical (integer / byte work: bit-exact), on calls in a row that share the
static array — a big mesh, then a small one that sees the big one's leftovers, then the big one again — and a triangle count
beyond the array's limit must be refused with an error that names the limit, leaving the buffers alone.

Collected last on purpose; independent of every other test: it passes alone in a fresh
process and after the rest of the suite."""
import numpy as np
import pytest

from rt_common import best_oracle, scenes, tracer_mod, vx

pytestmark = pytest.mark.gpu

MAX_TRIS = 65536   # include/vortex_b200.h: the static array's stated limit


def _device_for(mesh):
    sc = best_oracle().scene_create([mesh])
    karg = sc.setup(64, 48)
    bufs = sc.buffers()
    tr = tracer_mod.Tracer(64, 48, 1, 1).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    return tr, bufs


def _build(tr, bufs, n, static):
    """poison the two output buffers, build, download them whole"""
    for name in ("bvh", "triIdx"):
        tr.bufs[name].copy_to_dev(np.full(len(bufs[name]), 0xAB, np.uint8))
    fn = tr.dev.build_bvh_static if static else tr.dev.build_bvh
    assert fn(tr.bufs["tri"], n, tr.bufs["bvh"], tr.bufs["triIdx"]) == 2 * n - 1
    out = []
    for name in ("bvh", "triIdx"):
        a = np.zeros(len(bufs[name]), np.uint8)
        tr.bufs[name].copy_from_dev(a)
        out.append(a)
    return out


def test_static_build_equals_build_bvh():
    big, small = scenes.heightfield(64), scenes.icosphere((0, 0, 0), 1.0, 3)
    trs = {}
    try:
        for key, mesh in (("big", big), ("small", small)):
            tr, bufs = _device_for(mesh)
            n = mesh["tris"].shape[0]
            assert n <= MAX_TRIS
            trs[key] = (tr, bufs, n, _build(tr, bufs, n, static=False))
        # calls in a row on the shared array: every one must reproduce the allocating entry point's bytes
        for key in ("big", "small", "big", "big"):
            tr, bufs, n, want = trs[key]
            got = _build(tr, bufs, n, static=True)
            assert np.array_equal(got[1], want[1]), f"{key}: triIdx differs"
            assert np.array_equal(got[0], want[0]), f"{key}: bvh nodes differ"
            assert sorted(got[1].view(np.uint32)[:n].tolist()) == list(range(n))
            assert tr.dev.stats()["kernel_launches"] > 0
        # a frame over the tree the static build left behind is the frame over build_bvh's tree
        tr, bufs, n, _ = trs["big"]
        img_static = tr.run().copy()
        _build(tr, bufs, n, static=False)
        img = tr.run()
        assert np.array_equal(img, img_static) and len(np.unique(img)) > 4
    finally:
        for tr, *_ in trs.values():
            tr.close()


def test_static_build_refuses_sizes_beyond_the_array():
    mesh = scenes.icosphere((0, 0, 0), 1.0, 2)
    tr, bufs = _device_for(mesh)
    try:
        n = mesh["tris"].shape[0]
        before = _build(tr, bufs, n, static=True)
        with pytest.raises(vx.VxError) as ei:
            tr.dev.build_bvh_static(tr.bufs["tri"], MAX_TRIS + 1, tr.bufs["bvh"], tr.bufs["triIdx"])
        assert str(MAX_TRIS) in str(ei.value) and "limit" in str(ei.value)
        with pytest.raises(vx.VxError):
            tr.dev.build_bvh_static(tr.bufs["tri"], 0, tr.bufs["bvh"], tr.bufs["triIdx"])
        for name, want in zip(("bvh", "triIdx"), before):   # nothing was written by the refused calls
            a = np.zeros(len(bufs[name]), np.uint8)
            tr.bufs[name].copy_from_dev(a)
            assert np.array_equal(a, want)
        # and the entry point still works afterwards
        again = _build(tr, bufs, n, static=True)
        assert np.array_equal(again[0], before[0]) and np.array_equal(again[1], before[1])
    finally:
        tr.close()
