"""GPU parity tests (run on the B200 box): the CUDA path, called through the vx_* C ABI of libvortex-b200.so, against the CPU
oracle (the compiled reference oracle/_ref when it travelled with the snapshot, else the C port) and the golden fixtures.

Gates (BASELINE.json north_star): closest-hit IDs bit-exact on >= 99.99 % of rays with the residue attributed to equal-t
ties; t within 1e-5 relative; pixel colour within 1/255 per channel.  In strict order (VX_DCR_B200_RT_ORDER=1, the
reference's own visiting order) everything is required to be bit-identical."""
import glob
import os

import numpy as np
import pytest

from rt_common import ROOT, assert_hit_gates, best_oracle, channel_diff, hit_parity, pyoracle, scenes, tracer_for, tracer_mod, vx

pytestmark = pytest.mark.gpu
po = pyoracle()
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "ref_*.npz")))


@pytest.fixture(scope="module")
def oracle():
    return best_oracle()


def check_render(tr, sc, mode, strict, tol, kernel=0):
    tr.set_mode(mode, strict, kernel=kernel)
    img = tr.run()
    ref, _ = sc.render(mode=mode, threads=sc.o.hw_threads())
    d = channel_diff(img, ref)
    assert d <= tol, f"max channel diff {d} (mode={mode} strict={strict})"
    return img, ref


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
@pytest.mark.parametrize("strict", [1, 0])
def test_golden_fixtures(oracle, path, strict):
    """the reference's own outputs on the reference's own upload bytes"""
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    karg = g["karg"].view(po.KARG_DT).copy()
    hits_ref = g["hits"].view(po.HIT_DT)
    rs = po.RawScene(oracle, g, karg)
    tr = tracer_for(rs, karg)
    try:
        tr.set_mode(int(g["mode"]), strict)
        img = tr.run()
        d = channel_diff(img, g["image"])
        assert d <= (0 if strict else 1), d
        hp = hit_parity(tr.trace_primary(), hits_ref)
        assert_hit_gates(hp, strict)
        if strict:
            assert np.array_equal(img, g["image"])
    finally:
        tr.close()


CASES = [
    # name, meshes, w, h, depth, mode, setup kwargs
    ("cornell64_cfg0", lambda: [scenes.cornell_box()], 64, 64, 1, 0, {}),
    ("cornell_mirror", lambda: [scenes.cornell_box(0.5)], 160, 120, 4, 0, {}),
    ("hf224_shadow_cfg1", lambda: [scenes.heightfield(224)], 480, 270, 1, 1, {}),
    ("hf64_diffuse4_cfg2", lambda: [scenes.heightfield(64)], 192, 108, 4, 2, {}),
    ("multi_mirror", lambda: [scenes.heightfield(16), scenes.icosphere((0, 0.5, 0), 0.4, 2, reflectivity=0.5), scenes.cornell_box(0.3)],
     320, 240, 3, 0, {}),
    ("multi_shadow_diffuse", lambda: [scenes.heightfield(16), scenes.icosphere((0, 0.5, 0), 0.4, 2), scenes.cornell_box()],
     160, 120, 3, 3, {}),
    ("ragged_37x19", lambda: [scenes.heightfield(8)], 37, 19, 1, 0, {}),       # width/height not multiples of the 8x4 tile
    ("one_pixel", lambda: [scenes.cornell_box()], 1, 1, 1, 0, {}),
    ("spp3", lambda: [scenes.cornell_box(0.2)], 40, 24, 2, 0, {"spp": 3}),      # spp > 1 sums samples (no division), as the reference
]


@pytest.mark.parametrize("name,mk,w,h,depth,mode,kw", CASES, ids=[c[0] for c in CASES])
def test_render_and_hits_vs_oracle(oracle, name, mk, w, h, depth, mode, kw):
    sc = oracle.scene_create(mk())
    karg = sc.setup(w, h, max_depth=depth, **kw)
    tr = tracer_for(sc, karg)
    try:
        for kernel in (0, 1):     # 0 = ray state machine (default), 1 = one pixel per lane
            img_s, ref = check_render(tr, sc, mode, 1, 0, kernel)
            assert np.array_equal(img_s, ref)
            check_render(tr, sc, mode, 0, 1, kernel)
        tr.set_mode(mode, 0, kernel=0)
        ref_hits = sc.trace_primary(threads=oracle.hw_threads())
        for strict in (1, 0):
            tr.set_mode(mode, strict)
            assert_hit_gates(hit_parity(tr.trace_primary(), ref_hits), strict)
        st = tr.dev.stats()
        assert st["rays"] == w * h and st["kernel_launches"] >= 1
    finally:
        tr.close()


def test_incoherent_and_edge_rays(oracle):
    """caller-supplied rays: uniform random directions (config 5 shape, small) + axis-parallel / on-plane / grazing rays"""
    sc = oracle.scene_create([scenes.heightfield(96), scenes.cornell_box()])
    karg = sc.setup(16, 16, rotate_scene=False)
    rays = scenes.random_rays(200_000, (-1.5, -1.2, -1.5), (1.5, 1.2, 1.5), seed=5)
    o = np.array([[0, 0, -3], [-1, 0, -3], [1, 1, -3], [0, -1, -3], [0.25, 0.999, -3], [0, 0, 0], [-1, -1, -1], [0, 5, 0]], dtype=np.float32)
    d = np.array([[0, 0, 1], [0, 0, 1], [0, 0, 1], [0, 0, 1], [0, 0, 1], [0, 1, 0], [0.57735026, 0.57735026, 0.57735026], [0, -1, 0]], dtype=np.float32)
    rays = np.concatenate([rays, np.concatenate([o, d], axis=1)]).astype(np.float32)
    ref_hits = sc.trace_rays(rays, threads=oracle.hw_threads())
    tr = tracer_for(sc, karg)
    try:
        for strict in (1, 0):
            tr.set_mode(0, strict)
            assert_hit_gates(hit_parity(tr.trace_rays(rays), ref_hits), strict)
        assert tr.dev.stats()["rays"] == len(rays)
        # host-resident rays / records (vx_b200_trace_rays_host): same bytes as the buffer path for any chunking, ragged tail
        # included; the statistics are the sums over the chunks
        want = tr.trace_rays(rays)
        for chunk in ("4096", "65536", None):
            if chunk is None:
                os.environ.pop("VORTEX_B200_HOST_CHUNK", None)
            else:
                os.environ["VORTEX_B200_HOST_CHUNK"] = chunk
            got = tr.trace_rays_host(rays)
            assert got.tobytes() == want.tobytes(), chunk
            assert tr.dev.stats()["rays"] == len(rays)
        # coherence ordering of large batches (probe + octant/Morton sort + index array): same records at the same positions
        base_launches = tr.dev.stats()["kernel_launches"]
        os.environ["VORTEX_B200_RAY_ORDER_MIN"] = "1000"
        os.environ["VORTEX_B200_RAY_ORDER"] = "1"                 # opt-in (device-side ordering costs more than it gains)
        try:
            for bits in ("7", "3", "9"):
                os.environ["VORTEX_B200_RAY_ORDER_BITS"] = bits
                assert tr.trace_rays(rays).tobytes() == want.tobytes(), bits
            assert tr.dev.stats()["kernel_launches"] == 3            # keys + sort + trace: the random batch was reordered
            scan = np.zeros((4096, 6), np.float32)                   # a coherent batch (one scanline of parallel rays) is not
            scan[:, 0] = np.linspace(-1, 1, 4096)
            scan[:, 1], scan[:, 5] = 0.3, 1.0
            scan[:, 2] = -3.0
            tr.trace_rays(scan)
            assert tr.dev.stats()["kernel_launches"] == 2            # keys + trace
            os.environ["VORTEX_B200_RAY_ORDER"] = "0"
            assert tr.trace_rays(rays).tobytes() == want.tobytes()
            assert tr.dev.stats()["kernel_launches"] == 1
        finally:
            for k in ("VORTEX_B200_RAY_ORDER_MIN", "VORTEX_B200_RAY_ORDER_BITS", "VORTEX_B200_RAY_ORDER"):
                os.environ.pop(k, None)
        assert base_launches >= 1
        assert tr.trace_rays_host(rays[:1]).tobytes() == want[:1].tobytes()
        assert len(tr.trace_rays_host(rays[:0])) == 0
        import torch
        pr = torch.from_numpy(rays).pin_memory()
        ph = torch.empty(len(rays) * 24, dtype=torch.uint8).pin_memory()
        tr.dev.trace_rays_host(tr.args, pr.data_ptr(), len(rays), ph.data_ptr())
        assert ph.numpy().tobytes() == want.tobytes()
    finally:
        os.environ.pop("VORTEX_B200_HOST_CHUNK", None)
        tr.close()


def test_full_size_properties_config1():
    """BASELINE config 2 at its full size (1920x1080, 100 352 triangles, primary + shadow): size-independent properties.
    idempotence, strict == fast within the gates, and the union of a 2-way and a 3-way band partition == the whole image."""
    port = po.PortRaycast() if po.have_port() else best_oracle()
    sc = port.scene_create([scenes.heightfield(224)])
    karg = sc.setup(1920, 1080)
    tr = tracer_for(sc, karg)
    try:
        tr.set_mode(1, 0)
        a = tr.run().copy()
        b = tr.run().copy()
        assert np.array_equal(a, b)
        tr.set_mode(1, 1)
        s = tr.run().copy()
        assert channel_diff(a, s) <= 1
        assert (a != s).mean() <= 1e-4
        # a bounded sample of rows against the oracle (the whole frame is the bench's job)
        rows = (500, 508)
        ref, _ = sc.render(mode=1, rows=rows, threads=port.hw_threads())
        assert np.array_equal(s[rows[0]:rows[1]], ref[rows[0]:rows[1]])
        for world in (2, 3):
            acc = np.zeros_like(a)
            owner = (np.arange(1080) // 8) % world
            for rank in range(world):
                tr.dev.set_partition(rank, world, 8)
                tr.set_mode(1, 0)
                canvas = np.zeros_like(a)
                tr.out.copy_to_dev(canvas)
                part = tr.run().copy()
                assert not part[owner != rank].any()          # rows of other ranks untouched
                acc[owner == rank] = part[owner == rank]
            assert np.array_equal(acc, a)
        tr.dev.set_partition(0, 1, 8)
    finally:
        tr.close()


def test_stats_match_oracle_counters():
    """strict order visits exactly the nodes/triangles the reference visits: the GPU work counters equal the port's"""
    if not po.have_port():
        pytest.skip("C port not built")
    port = po.PortRaycast()
    sc = port.scene_create([scenes.heightfield(32), scenes.icosphere((0, 0.4, 0), 0.3, 1, reflectivity=0.5)])
    karg = sc.setup(128, 96, max_depth=3)
    _, cnt = sc.render_counted(mode=0, threads=2)
    tr = tracer_for(sc, karg)
    try:
        tr.set_mode(0, 1)
        tr.dev.enable_stats(True)
        tr.run()
        st = tr.dev.stats()
        for k in ("rays", "box_tests", "tri_tests", "blas_entered", "shaded_hits"):
            assert st[k] == cnt[k], (k, st[k], cnt[k])
        # RT-unit style counters (RTUnit::PerfStats): SIMT efficiency of the run is a fraction, and without statistics they are 0
        eff = st["lane_steps"] / (32.0 * st["warp_steps"])
        assert st["warps"] > 0 and 0.0 < eff <= 1.0
        tr.dev.enable_stats(False)
        tr.run()
        st = tr.dev.stats()
        assert st["rays"] == cnt["rays"] and st["warp_steps"] == 0 and st["lane_steps"] == 0
    finally:
        tr.close()


def test_abi_error_behaviour():
    """argument checks of the glue layer (reference: runtime/common/callbacks.inc) — non-zero on misuse, 0 on null free"""
    import ctypes as C
    L = vx.lib()
    dev = vx.Device()
    try:
        b = C.c_void_p()
        assert L.vx_mem_alloc(dev.h, 0, 3, C.byref(b)) != 0                 # size 0 rejected
        assert L.vx_mem_alloc(None, 64, 3, C.byref(b)) != 0                 # null device
        assert L.vx_mem_free(None) == 0                                     # null buffer frees fine
        buf = dev.mem_alloc(256)
        host = np.zeros(512, dtype=np.uint8)
        assert L.vx_copy_to_dev(buf.h, host.ctypes.data, 0, 512) != 0       # beyond the buffer
        assert L.vx_copy_to_dev(buf.h, host.ctypes.data, 200, 100) != 0     # offset + size beyond
        assert L.vx_copy_from_dev(host.ctypes.data, buf.h, 0, 512) != 0
        assert L.vx_copy_to_dev(buf.h, None, 0, 16) != 0                    # null host pointer
        src = np.arange(256, dtype=np.uint8)
        buf.copy_to_dev(src)
        back = np.zeros(256, dtype=np.uint8)
        buf.copy_from_dev(back)
        assert np.array_equal(src, back)
        a = buf.address
        assert a >= 0x10000 and a < (1 << 32) and a % 64 == 0               # simx-style address space
        free0, used0 = dev.mem_info()
        buf.free()
        free1, used1 = dev.mem_info()
        assert used1 < used0 and free1 > free0
        k = dev.upload_kernel_bytes(vx.DUMMY_KERNEL_IMAGE)                  # reserve at 0x80000000
        assert k.address == 0x80000000
        with pytest.raises(vx.VxError):
            dev.upload_kernel_bytes(vx.DUMMY_KERNEL_IMAGE)                  # same range twice
        args = dev.upload_bytes(np.zeros(16, dtype=np.uint8))
        assert L.vx_start(dev.h, k.h, args.h) != 0                          # 16-byte argument block: neither layout
        assert b"kernel_arg_t" in L.vx_b200_last_error()
        assert dev.caps(0x3) >= 100 and dev.caps(0x1) == 32                 # NUM_CORES = SMs, NUM_THREADS = warp size
        v = C.c_uint64()
        assert L.vx_dev_caps(dev.h, 0x99, C.byref(v)) != 0                  # unknown caps id: error, not abort
        dev.dcr_write(0x001, 0x1234)
        assert dev.dcr_read(0x001) == 0x1234
        assert dev.ready_wait(10) is None                                   # nothing pending
    finally:
        dev.close()


REF_APP = os.path.join(ROOT, "oracle", "_ref", "raycast")


@pytest.mark.skipif(not (os.path.exists(REF_APP) and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "assets", "teapot.obj"))),
                    reason="the compiled reference app (oracle/_ref/raycast + assets) did not travel with the snapshot")
@pytest.mark.parametrize("flags,md5", [(["-w64", "-h64"], "a8248b0896d09d1b3b98b67f5c43ce6c"),
                                       (["-n3", "-d4", "-w64", "-h64"], "d8ecb7285f1920568d8f32abe0ad600c")],
                         ids=["teapot", "three_teapots_mirror_d4"])
@pytest.mark.parametrize("order", ["1", "0"], ids=["strict", "fast"])
def test_unmodified_reference_app_runs_on_b200(tmp_path, flags, md5, order):
    """drop-in proof: the reference's own raycast binary (its Tracer, Scene, BVH, OBJ/PNG loaders, linked against the reference
    stub libvortex.so) with VORTEX_DRIVER=b200 -> dlopen(libvortex-b200.so) -> vx_dev_init.  The PPM it writes must be
    byte-identical to the one its own CPU path (-c) writes (md5 recorded in BASELINE.md 2)."""
    import hashlib
    import subprocess
    lib_dir = os.path.dirname(vx.LIB_PATH)
    kernel = tmp_path / "kernel.vxbin"
    kernel.write_bytes(vx.DUMMY_KERNEL_IMAGE)
    out = tmp_path / "out.ppm"
    env = dict(os.environ, VORTEX_DRIVER="b200", VORTEX_B200_ORDER=order,
               LD_LIBRARY_PATH=lib_dir + ":" + os.path.dirname(REF_APP) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([REF_APP, "-k", str(kernel), "-o", str(out)] + flags, cwd=os.path.dirname(REF_APP), env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout
    assert hashlib.md5(out.read_bytes()).hexdigest() == md5, r.stdout[-400:]


def test_schedule_feedback_keeps_the_frame():
    """frame-to-frame schedule feedback: frames 0-1 run row-major, frame 2 in descending-cost order, later frames in
    whichever order won the trial — the picture must be the same in every phase, for the whole image and for a band partition,
    and equal to a run with the feedback disabled"""
    port = po.PortRaycast() if po.have_port() else best_oracle()
    sc = port.scene_create([scenes.heightfield(96), scenes.icosphere((0, 0.4, 0), 0.3, 2, reflectivity=0.5)])
    karg = sc.setup(960, 540, max_depth=2)
    os.environ["VORTEX_B200_SCHEDULE"] = "0"
    try:
        tr = tracer_for(sc, karg)
        try:
            tr.set_mode(1, 0)
            want = tr.run().copy()
        finally:
            tr.close()
    finally:
        os.environ.pop("VORTEX_B200_SCHEDULE", None)
    tr = tracer_for(sc, karg)
    try:
        for mode in (1, 0):
            tr.set_mode(mode, 0)
            ref = want if mode == 1 else None
            for frame in range(6):
                got = tr.run()
                if ref is None:
                    ref = got.copy()
                assert np.array_equal(got, ref), (mode, frame)
        tr.set_mode(1, 0)
        parts = np.zeros_like(want)
        for rank in range(2):
            tr.dev.set_partition(rank, 2, 8)
            for frame in range(5):
                got = tr.run()
            rows = np.concatenate([np.arange(b * 8, min(b * 8 + 8, 540)) for b in range(rank, (540 + 7) // 8, 2)])
            parts[rows] = got[rows]
        assert np.array_equal(parts, want)
    finally:
        tr.close()


@pytest.mark.gpu
def test_gpu_gather_probe_reports_a_bandwidth():
    """vx_b200_probe_gather (roofline denominator): a plausible GB/s figure for 32- and 128-byte records (on B200 the 32-byte
    gather measures ~6.0 TB/s in L2 and the 128-byte one ~4.6 TB/s, so no ordering is asserted), bad record sizes are refused"""
    dev = vx.Device()
    try:
        g32 = dev.probe_gather(32 << 20, 32)
        g128 = dev.probe_gather(32 << 20, 128)
        assert 100.0 < g32 < 50000.0 and 100.0 < g128 < 50000.0, (g32, g128)
        with pytest.raises(vx.VxError):
            dev.probe_gather(32 << 20, 48)
    finally:
        dev.close()


def test_partial_upload_into_large_instance_buffer(oracle):
    """a host that patches ONE blas_node_t of a > 4 KB instance buffer (27 instances x 160 B) with a small vx_copy_to_dev: the
    backend's host copy of small uploads must not be mistaken for the whole buffer (instance bases of the 4-wide records)"""
    # 27 instances of ONE mesh placed by their transforms, as the reference app does with -n (raycast/tracer.cpp:89-100).  (Meshes
    # that differ are not usable here: the reference's BVH leaves index triIdx relative to the mesh's own slice while render.h
    # indexes the global buffer (raycast/scene.cpp:62 vs render.h:88-90), so instance i > 0 shows mesh 0's triangles.)
    def shift(x, y, z):
        m = np.eye(4, dtype=np.float32)
        m[0, 3], m[1, 3], m[2, 3] = x, y, z
        return m
    base = scenes.icosphere((0, 0, 0), 0.17, 1)
    meshes = [dict(base, transform=shift(0.4 * (i % 3 - 1), 0.4 * ((i // 3) % 3 - 1), 0.4 * (i // 9 - 1))) for i in range(27)]
    sc = oracle.scene_create(meshes)
    karg = sc.setup(320, 240)
    ref, _ = sc.render(mode=0, threads=oracle.hw_threads())
    tr = tracer_for(sc, karg)
    try:
        blas = tr.host_scene["blas"]
        assert blas.nbytes == 27 * 160 > 4096
        for strict in (1, 0):
            tr.set_mode(0, strict)
            assert channel_diff(tr.run(), ref) <= (0 if strict else 1)
            for k in (5, 26):
                tr.bufs["blas"].copy_to_dev(np.ascontiguousarray(blas[k * 160:(k + 1) * 160]), dst_offset=k * 160)
                assert channel_diff(tr.run(), ref) <= (0 if strict else 1), (strict, k)
        # moving one instance through a small upload is seen by the next frame
        moved = blas.copy().view(np.float32).reshape(27, 40)
        moved[3, 16 + 3] += 0.1                      # invTransform translation x of instance 3 (cells 3 / 7 / 11)
        tr.bufs["blas"].copy_to_dev(np.ascontiguousarray(moved[3]).view(np.uint8), dst_offset=3 * 160)
        img = tr.run()
        assert (img != ref).any()
        rs = po.RawScene(oracle, dict(tr.host_scene, blas=moved.view(np.uint8).reshape(-1)), karg)
        ref2, _ = rs.render(mode=0, threads=oracle.hw_threads())
        assert channel_diff(img, ref2) <= 1
    finally:
        tr.close()


def test_argument_buffer_with_headroom_is_a_raycast_launch(oracle):
    """a raycast host whose kernel_arg_t sits in a larger buffer (256 bytes, or at an offset) is still layout A: the layout is
    taken from the RT-unit DCRs / VX_DCR_B200_RT_LAYOUT, the size is only a lower bound"""
    sc = oracle.scene_create([scenes.cornell_box()])
    karg = sc.setup(96, 64)
    tr = tracer_for(sc, karg)
    try:
        tr.set_mode(0, 1)
        want = tr.run().copy()
        big = tr.dev.mem_alloc(256)
        big.copy_to_dev(np.frombuffer(tr.karg.tobytes() + b"\0" * 64, dtype=np.uint8))
        tr.out.copy_to_dev(np.zeros((64, 96), np.uint32))
        tr.dev.start(tr.krnl, big)
        tr.dev.ready_wait()
        got = np.zeros((64, 96), np.uint32)
        tr.out.copy_from_dev(got)
        assert np.array_equal(got, want)
        big.free()
    finally:
        tr.close()


def test_deep_tree_overflow_is_reported_not_silent():
    """a degenerate BVH2 deeper than the reference's own stack[64] (render.h:80): every push is bounds-checked, the run is
    reported as failed (error bit 0) instead of writing past the traversal stack"""
    n = 200                                           # a chain: every inner node = (next inner node, one leaf); all boxes equal
    tris = np.zeros((n, 9), np.float32)
    tris[:, 0:3] = (-1, -1, 0)
    tris[:, 3:6] = (1, -1, 0)
    tris[:, 6:9] = (0, 1, 0)
    tris[:, 2::3] += (np.arange(n, dtype=np.float32) * 0.01)[:, None]
    nodes = np.zeros(2 * n, dtype=[("mn", "<f4", 3), ("leftFirst", "<u4"), ("mx", "<f4", 3), ("triCount", "<u4")])
    nodes["mn"], nodes["mx"] = (-1, -1, -1), (1, 1, 3)
    # inner k (root = node 0, else node 2k-1) has children 2k+1 (the next inner node; equal distances => visited first in strict
    # order) and 2k+2 (leaf k, pushed): one stack entry per level
    for k in range(n - 1):
        inner = 0 if k == 0 else 2 * k - 1
        nodes[inner]["leftFirst"], nodes[inner]["triCount"] = 2 * k + 1, 0
        nodes[2 * k + 2]["leftFirst"], nodes[2 * k + 2]["triCount"] = k, 1
    nodes[2 * n - 3]["leftFirst"], nodes[2 * n - 3]["triCount"] = n - 1, 1
    port = po.PortRaycast() if po.have_port() else best_oracle()
    sc = port.scene_create([{"tris": tris, "triex": np.zeros((n, 15), np.float32), "tex": np.full((2, 2), 0x808080, np.uint32)}])
    karg = sc.setup(32, 32)
    bufs = dict(sc.buffers())
    bufs["bvh"] = nodes.view(np.uint8).reshape(-1).copy()
    bufs["triIdx"] = np.arange(n, dtype=np.uint32).view(np.uint8)
    tr = tracer_mod.Tracer(32, 32, 1, 1).init(bufs)
    try:
        tr.setup({k: karg[k][0] for k in karg.dtype.names})
        for kernel in (0, 1):
            tr.set_mode(0, 1, kernel=kernel)          # strict order pushes one entry per level: depth 200 > 128 / 64
            with pytest.raises(vx.VxError, match="stack overflow"):
                tr.run()
    finally:
        tr.close()


def test_readback_strips_deliver_the_same_frame():
    """vx_b200_set_readback: the frame is in page-locked host memory when vx_ready_wait returns.  At benchmark sizes the frame is
    cut into strips (two kernel streams + a copy stream); any strip count must give the image of the plain call sequence."""
    import torch
    port = po.PortRaycast() if po.have_port() else best_oracle()
    sc = port.scene_create([scenes.heightfield(224)])
    w, h = 2048, 1100                                   # ragged last band (1100 = 137 bands + 4 rows)
    karg = sc.setup(w, h)
    tr = tracer_for(sc, karg)
    os.environ["VORTEX_B200_SCHEDULE"] = "0"          # row-major tile order on every frame (an ordered schedule runs as one launch)
    os.environ["VORTEX_B200_READBACK_MIN_TILES"] = "1000"   # strips normally start at 8K-sized shares (200 000 tiles)
    try:
        for mode in (1, 0):
            tr.set_mode(mode, 0)
            want = tr.run().copy()
            host = torch.zeros((h, w), dtype=torch.int32).pin_memory()
            tr.dev.set_readback(tr.out, host.data_ptr(), w * h * 4)
            for strips in ("4", "1", "7", "3"):
                os.environ["VORTEX_B200_READBACK_STRIPS"] = strips
                for frame in range(6):                  # frames 0-2 run the schedule trial (no strips), later ones the strips
                    host.zero_()
                    tr.dev.start(tr.krnl, tr.args)
                    tr.dev.ready_wait()
                    assert np.array_equal(host.numpy().view(np.uint32), want), (mode, strips, frame)
                    got = np.zeros((h, w), np.uint32)
                    tr.out.copy_from_dev(got)           # the device framebuffer is complete as well
                    assert np.array_equal(got, want)
                    st = tr.dev.stats()
                    assert st["kernel_launches"] - st["prep_launches"] == int(strips), (strips, st)   # the strips did run
            tr.dev.set_readback(tr.out, None, 0)
    finally:
        os.environ.pop("VORTEX_B200_READBACK_STRIPS", None)
        os.environ.pop("VORTEX_B200_SCHEDULE", None)
        os.environ.pop("VORTEX_B200_READBACK_MIN_TILES", None)
        tr.close()


def test_readback_strip_order_with_a_band_partition():
    """The strips may go top-down or bottom-up (VORTEX_B200_READBACK_ORDER; by default the previous frame's timing decides) and may
    be tapered; with a band partition (what every GPU of a group has) a rank's strips must deliver exactly that rank's bands —
    including the short last band — and touch nothing else of the host image."""
    import torch
    port = po.PortRaycast() if po.have_port() else best_oracle()
    sc = port.scene_create([scenes.heightfield(224)])
    w, h = 2048, 1100                                   # 137 bands of 8 rows + 4 rows
    karg = sc.setup(w, h)
    tr = tracer_for(sc, karg)
    os.environ["VORTEX_B200_SCHEDULE"] = "0"
    os.environ["VORTEX_B200_READBACK_MIN_TILES"] = "1000"
    keys = ("VORTEX_B200_READBACK_STRIPS", "VORTEX_B200_READBACK_TAPER", "VORTEX_B200_READBACK_ORDER")
    try:
        tr.set_mode(1, 0)
        want = tr.run().copy()
        host = torch.zeros((h, w), dtype=torch.int32).pin_memory()
        tr.dev.set_readback(tr.out, host.data_ptr(), w * h * 4)
        band_of_row = np.arange(h) // 8
        for world in (2, 3, 1):
            for rank in range(world):
                tr.dev.set_partition(rank, world, 8)
                mine = (band_of_row % world) == rank
                for strips, taper, order in (("4", "100", "1"), ("5", "80", "1"), ("5", "80", "0"), ("4", "100", "2")):
                    for k, v in zip(keys, (strips, taper, order)):
                        os.environ[k] = v
                    for frame in range(3 if order == "2" else 1):
                        host.zero_()
                        tr.dev.start(tr.krnl, tr.args)
                        tr.dev.ready_wait()
                        got = host.numpy().view(np.uint32)
                        assert np.array_equal(got[mine], want[mine]), (world, rank, strips, taper, order, frame)
                        assert not got[~mine].any(), (world, rank, strips, taper, order, frame)
                        st = tr.dev.stats()
                        assert st["kernel_launches"] - st["prep_launches"] == int(strips), (strips, st)
        tr.dev.set_partition(0, 1, 8)
        tr.dev.set_readback(tr.out, None, 0)
    finally:
        for k in keys + ("VORTEX_B200_SCHEDULE", "VORTEX_B200_READBACK_MIN_TILES"):
            os.environ.pop(k, None)
        tr.close()
