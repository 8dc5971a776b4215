"""GPU BVH build (vx_b200_build_bvh): the tree it leaves in the Vortex buffers is the reference's layout-A byte
format, so the checks run the REFERENCE's own traversal (CPU oracle, RawScene) on the downloaded buffers:
  * structure: 2N-1 nodes, children adjacent, every triangle in exactly one leaf, boxes contain their subtree;
  * the oracle on the GPU-built tree finds the same closest-hit distances as on the host-built (reference-identical) tree;
  * the CUDA kernels on the GPU-built tree equal the oracle on the same bytes (strict order bit for bit)."""
import numpy as np
import pytest

from rt_common import assert_hit_gates, best_oracle, hit_parity, pyoracle, scenes, tracer_mod, vx

pytestmark = pytest.mark.gpu
po = pyoracle()
BVH_DT = np.dtype([("lo", "<f4", 3), ("leftFirst", "<u4"), ("hi", "<f4", 3), ("triCount", "<u4")])


def gpu_built_scene(oracle, mesh, w, h, leaf_max=None):
    sc = oracle.scene_create([mesh])
    karg = sc.setup(w, h)
    bufs = sc.buffers()
    n = mesh["tris"].shape[0]
    tr = tracer_mod.Tracer(w, h, 1, 1).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    import os
    if leaf_max is not None:
        os.environ["VORTEX_B200_BUILD_MAX_LEAF"] = str(leaf_max)
    try:
        used = tr.dev.build_bvh(tr.bufs["tri"], n, tr.bufs["bvh"], tr.bufs["triIdx"])
    finally:
        os.environ.pop("VORTEX_B200_BUILD_MAX_LEAF", None)
    assert used == 2 * n - 1
    st = tr.dev.stats()
    bvh = np.zeros(len(bufs["bvh"]), np.uint8)
    idx = np.zeros(len(bufs["triIdx"]), np.uint8)
    tr.bufs["bvh"].copy_from_dev(bvh)
    tr.bufs["triIdx"].copy_from_dev(idx)
    gb = dict(bufs)
    gb["bvh"], gb["triIdx"] = bvh, idx
    # TLAS leaf box = the root box, which any correct build reproduces exactly
    nodes = bvh.view(BVH_DT)
    host_root = bufs["bvh"].view(BVH_DT)[0]
    assert np.array_equal(nodes[0]["lo"], host_root["lo"]) and np.array_equal(nodes[0]["hi"], host_root["hi"])
    return sc, karg, bufs, gb, tr, st


@pytest.mark.parametrize("name,mk", [("cornell", lambda: scenes.cornell_box()), ("hf64", lambda: scenes.heightfield(64)),
                                     ("ico3", lambda: scenes.icosphere((0, 0, 0), 1.0, 3))], ids=["cornell", "hf64", "ico3"])
@pytest.mark.parametrize("leaf_max", [4, 1, 14], ids=["leaf4_default", "leaf1", "leaf14"])
def test_gpu_built_bvh(name, mk, leaf_max):
    oracle = best_oracle()
    mesh = mk()
    n = mesh["tris"].shape[0]
    sc, karg, bufs, gb, tr, st = gpu_built_scene(oracle, mesh, 96, 72, leaf_max)
    try:
        nodes = gb["bvh"].view(BVH_DT)[: 2 * n - 1]
        idx = gb["triIdx"].view(np.uint32)
        assert sorted(idx.tolist()) == list(range(n))
        # the tree reachable from the root (nodes below a subtree that was priced into one leaf stay behind as fillers)
        reach, todo = [], [0]
        while todo:
            k = todo.pop()
            reach.append(k)
            if nodes["triCount"][k] == 0:
                todo += [int(nodes["leftFirst"][k]), int(nodes["leftFirst"][k]) + 1]
        reach = np.array(sorted(reach))
        assert len(set(reach.tolist())) == len(reach)                       # every node referenced once
        rn = nodes[reach]
        leaves = rn[rn["triCount"] != 0]
        slots = np.concatenate([np.arange(f, f + c) for f, c in zip(leaves["leftFirst"], leaves["triCount"])])
        assert sorted(slots.tolist()) == list(range(n))                     # every triangle in exactly one leaf
        assert leaves["triCount"].max() <= leaf_max
        inner = reach[nodes["triCount"][reach] == 0]
        assert len(inner) == len(leaves) - 1
        # boxes: a parent's box is the union of its two (adjacent) children; a leaf's box is its triangles' vertex min/max
        l, r = nodes[nodes["leftFirst"][inner]], nodes[nodes["leftFirst"][inner] + 1]
        assert np.array_equal(nodes["lo"][inner], np.minimum(l["lo"], r["lo"])) and np.array_equal(nodes["hi"][inner], np.maximum(l["hi"], r["hi"]))
        tri = bufs["tri"].view(np.float32).reshape(-1, 3, 3)
        for lf in leaves[:: max(1, len(leaves) // 500)]:
            v = tri[idx[lf["leftFirst"]: lf["leftFirst"] + lf["triCount"]]].reshape(-1, 3)
            assert np.array_equal(lf["lo"], v.min(axis=0)) and np.array_equal(lf["hi"], v.max(axis=0))
        if leaf_max == 1:
            assert len(leaves) == n
        elif name == "hf64":
            assert len(leaves) < 0.75 * n                                   # the two triangles of a quad share a box: one leaf
        # the reference's traversal on the GPU-built tree vs on the host-built tree
        raw = po.RawScene(oracle, gb, karg)
        h_gpu_tree = raw.trace_primary(threads=oracle.hw_threads())
        h_host_tree = sc.trace_primary(threads=oracle.hw_threads())
        assert np.array_equal(h_gpu_tree["dist"], h_host_tree["dist"])
        assert_hit_gates(hit_parity(h_gpu_tree, h_host_tree), False)
        # the CUDA kernels on the GPU-built tree vs the oracle on the same bytes
        for strict in (1, 0):
            tr.set_mode(0, strict)
            assert_hit_gates(hit_parity(tr.trace_primary(), h_gpu_tree), strict)
        img = tr.run()
        ref_img, _ = raw.render(threads=oracle.hw_threads())
        assert np.abs(((img >> 8) & 255).astype(int) - ((ref_img >> 8) & 255).astype(int)).max() <= 1
        assert st["kernel_launches"] >= 5
    finally:
        tr.close()


def test_gpu_build_one_triangle_and_errors():
    one = scenes.heightfield(1)
    mesh = {"tris": one["tris"][:1], "triex": one["triex"][:1], "tex": one["tex"], "reflectivity": 0.0}
    oracle = best_oracle()
    sc, karg, bufs, gb, tr, st = gpu_built_scene(oracle, mesh, 8, 8)
    try:
        root = gb["bvh"].view(BVH_DT)[0]
        assert root["triCount"] == 1 and root["leftFirst"] == 0
        with pytest.raises(vx.VxError):
            tr.dev.build_bvh(tr.bufs["tri"], 5, tr.bufs["bvh"], tr.bufs["triIdx"])   # more triangles than the buffer holds
    finally:
        tr.close()


def test_gpu_qbvh_build_layout_b():
    """vx_b200_build_qbvh (layout B): the GPU-built quantized BVH4 + leaf-ordered triangles, downloaded, are valid
    RT-unit input — every triangle in exactly one leaf, child boxes conservative under the RT unit's decode — and the reference's
    traversal (port / compiled reference, through check_scene) finds on them what the GPU finds; distances equal those on the
    host-built tree of the same mesh."""
    import importlib
    import os
    from rtu_common import b_mesh, rtu, tracer_b
    from test_gpu_rtu import check_scene
    R = rtu()
    host = importlib.import_module("vortex-raytracing_b200.host")
    mesh = b_mesh(scenes.heightfield(40), two_materials=True)
    n = mesh["tris"].shape[0]
    hs = host.RtxHostScene([mesh])
    karg = hs.setup(128, 96)
    bufs = hs.buffers()
    tr = tracer_b(bufs, karg, 1)
    try:
        hits_host = tr.trace_primary().copy()
        used, box = tr.dev.build_qbvh(tr.bufs["tri"], n, tr.bufs["triEx"], tr.bufs["qbvh"])
        assert n < used <= 2 * n
        built = dict(bufs)
        for name in ("tri", "triEx", "qbvh"):
            arr = np.empty_like(bufs[name])
            tr.bufs[name].copy_from_dev(arr)
            built[name] = arr
        built["bvh_nodes_used"] = used
        hits_gpu = tr.trace_primary().copy()          # the rebuilt scene, straight from device memory
    finally:
        tr.close()
    src = mesh["tris"].reshape(-1, 9).astype(np.float32)
    out = built["tri"].view(np.float32).reshape(-1, 9)
    assert sorted(map(bytes, src)) == sorted(map(bytes, out))                      # a permutation of the input
    ex_src = bufs["triEx"].view(np.uint32).reshape(-1, 16)
    ex_out = built["triEx"].view(np.uint32).reshape(-1, 16)
    key = {bytes(t): bytes(e) for t, e in zip(bufs["tri"].view(np.float32).reshape(-1, 9), ex_src)}
    assert all(key[bytes(t)] == bytes(e) for t, e in zip(out, ex_out))              # attributes moved with their triangles
    assert np.array_equal(box[:3], src.reshape(-1, 3).min(axis=0)) and np.array_equal(box[3:], src.reshape(-1, 3).max(axis=0))
    q = built["qbvh"].view(R.QNODE_DT)[:used]
    assert (q["imask"] == 0).all()
    leaves = q[q["leafIdx"] != 0]
    assert len(leaves) == n and sorted(int(x) for x in leaves["leftFirst"]) == list(range(n))
    tri = out.reshape(-1, 3, 3)
    inner = np.nonzero(q["leafIdx"] == 0)[0]
    seen = np.zeros(used, dtype=np.int32)
    seen[0] = 1
    for i in inner:                                                                 # every node: conservative child boxes
        nd = q[i]
        e = np.array([nd["ex"], nd["ey"], nd["ez"]], dtype=np.int32)
        for k in range(4):
            c = nd["children"][k]
            if not c["meta"]:
                continue
            ci = int(nd["leftFirst"]) + k
            seen[ci] += 1
            lo = nd["origin"] + np.ldexp(c["q"][:3].astype(np.float32), e)
            hi = nd["origin"] + np.ldexp(c["q"][3:].astype(np.float32), e)
            child = q[ci]
            if child["leafIdx"] != 0:
                v = tri[child["leftFirst"]].reshape(-1, 3)
                assert (v >= lo).all() and (v <= hi).all()
            else:                                                                   # an inner child's own origin lies inside
                assert (child["origin"] >= lo).all() and (child["origin"] <= hi).all()
    assert (seen == 1).all()                                                        # a tree: every node referenced once
    # same geometry, same arithmetic: the distances found on the GPU-built tree equal those on the host-built one
    assert np.array_equal(hits_gpu["dist"], hits_host["dist"])
    assert (hits_gpu["dist"] < 1e30).mean() > 0.1
    check_scene(built, karg, 1)
