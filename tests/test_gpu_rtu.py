"""GPU parity tests of the RT-unit (layout B) path: the CUDA kernels, called through the vx_* C ABI of libvortex-b200.so
with the reference raytracing host's call sequence (RtxTracer), against the CPU oracles on the SAME uploaded bytes.

Oracles: oracle/_ref/libref_rtu.so (sim/simx/rt_traversal.cpp compiled verbatim) when it travelled with the snapshot, and
the plain-C port (bit-identical to it, tests/test_oracle_rtu.py).  The compiled reference spins for 2^32 iterations per
occurrence of its `size()-1` wrap (rt_traversal.cpp:81), so rays that reach it (flagged by the port) are compared with the
port only.

Gates (BASELINE.json north_star): closest-hit IDs bit-exact on >= 99.99 % of rays, residue = equal-t ties; t within 1e-5
relative; colour within 1/255.  In strict order (VX_DCR_B200_RT_ORDER=1: the reference's trail/short-stack algorithm visit
for visit) hit records, frames AND work counters are required to be identical."""
import glob
import os

import numpy as np
import pytest

from rt_common import ROOT, assert_hit_gates, channel_diff, hit_parity
from rtu_common import SCENES_B, rtu, tracer_b

pytestmark = pytest.mark.gpu
R = rtu()
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "rtu_*.npz")))


def load_golden(path):
    z = np.load(path)
    g = {k: z[k] for k in z.files}
    karg = g["karg"].view(R.KARG_B_DT).copy()
    return g, karg


def check_scene(bufs, karg, cam, ref_hits=None, ref_image=None):
    port = R.RtuPort()
    img = R.Image(bufs, karg)
    threads = os.cpu_count() or 1
    rays = port.primary_rays(img, cam)
    hits_port, work, flags = port.trace_rays(img, rays, want_work=True, want_wrap_flags=True)
    frame_port, work_r = port.render(img, cam, threads=threads, want_work=True)
    # The reference addresses the children of a popped TLAS inner node from a stale base_ptr (rt_traversal.cpp:179-213) and so
    # misses geometry behind TLASes of more than one inner level.  Strict order reproduces that address for address; the fast
    # order does not have the defect and is checked against the port with the base fixed (oracle/port/rtu_port.c).
    port.set_fix_tlas_base(1)
    try:
        hits_fix = port.trace_rays(img, rays, threads=threads)
        frame_fix = port.render(img, cam, threads=threads)
    finally:
        port.set_fix_tlas_base(0)
    tr = tracer_b(bufs, karg, cam)
    try:
        tr.dev.enable_stats(True)
        for strict in (1, 0):
            tr.set_order(strict)
            want_hits, want_frame = (hits_port, frame_port) if strict else (hits_fix, frame_fix)
            hits = tr.trace_primary()
            st = tr.dev.stats()
            assert_hit_gates(hit_parity(hits, want_hits), strict)
            if strict:
                assert hits.tobytes() == hits_port.tobytes()
                # work counters of the reference algorithm: node reads, instance entries, triangle tests
                assert (st["box_tests"], st["blas_entered"], st["tri_tests"]) == (work["nodes"], work["blas_entered"], work["tri_tests"]), (st, work)
                # RT-unit style counters: one traversal step per node the reference reads
                assert st["lane_steps"] == work["nodes"]
            assert st["warps"] > 0 and 0 < st["lane_steps"] <= 32 * st["warp_steps"]
            # caller-supplied rays take the same path, from a device buffer or straight from host memory in chunks
            assert tr.trace_rays(rays).tobytes() == hits.tobytes()
            os.environ["VORTEX_B200_HOST_CHUNK"] = "4096"
            try:
                assert tr.trace_rays_host(rays).tobytes() == hits.tobytes()
            finally:
                os.environ.pop("VORTEX_B200_HOST_CHUNK", None)
            frame = tr.run()
            d = channel_diff(frame, want_frame)
            assert d <= (0 if strict else 1), (strict, d, int((frame != want_frame).sum()))
            if strict:
                assert np.array_equal(frame, frame_port)
                assert tr.dev.stats()["shaded_hits"] == work_r["shaded_hits"]
                if ref_hits is not None:   # outputs of the compiled reference recorded in the fixture
                    assert hits.tobytes() == ref_hits.tobytes()
                    assert np.array_equal(frame, ref_image)
            fused = tr.dev.stats()                      # of the frame above
            if not strict:
                # fast-order hit records come from the warp-scheduled ray state machine (rt_machine_b.cu), frames from the
                # one-ray-per-lane kernel (rt_rtu.cu) unless VORTEX_B200_RTU_MACHINE=2: same records, same frame, same work
                fused_1 = fused
                os.environ["VORTEX_B200_RTU_MACHINE"] = "2"
                try:
                    assert tr.trace_primary().tobytes() == hits.tobytes()
                    assert np.array_equal(tr.run(), frame)
                    mach = tr.dev.stats()
                    for k in ("rays", "box_tests", "tri_tests", "blas_entered", "shaded_hits"):
                        assert mach[k] == fused_1[k], (k, mach[k], fused_1[k])
                    os.environ["VORTEX_B200_RTU_MACHINE"] = "0"
                    assert tr.trace_primary().tobytes() == hits.tobytes()
                    os.environ["VORTEX_B200_RTU_DIRECT"] = "0"      # rt_rtu.cu's k_rtu_render (nodes without child cursors)
                    assert np.array_equal(tr.run(), frame)
                    old = tr.dev.stats()
                    for k in ("rays", "box_tests", "tri_tests", "blas_entered", "shaded_hits"):
                        assert old[k] == fused_1[k], (k, old[k], fused_1[k])
                finally:
                    os.environ.pop("VORTEX_B200_RTU_MACHINE", None)
                    os.environ.pop("VORTEX_B200_RTU_DIRECT", None)
            # the three shipped shaders restated against the traceRay/getWork/getAttr/commit shim and dispatched through the
            # SBT: the same frame and the same work as the fused kernel, in either order
            tr.set_program(0)
            try:
                assert np.array_equal(tr.run(), frame)
                prog = tr.dev.stats()
                for k in ("rays", "box_tests", "tri_tests", "blas_entered", "shaded_hits"):
                    assert prog[k] == fused[k], (k, strict, prog[k], fused[k])
            finally:
                tr.set_program(None)
        if R.have_ref():
            ok = flags == 0
            h_ref = R.RtuRef().trace_rays(img, rays[ok], threads=threads)
            tr.set_order(1)
            assert tr.trace_rays(rays[ok]).tobytes() == h_ref.tobytes()
    finally:
        tr.close()


@pytest.mark.parametrize("name", list(SCENES_B), ids=list(SCENES_B))
def test_rtu_scene_vs_oracle(name):
    mk, w, h, depth, cam = SCENES_B[name]
    if R.have_ref():      # the reference's own BVH4 builder + quantizer
        bufs, karg = R.RtuRef().build_scene(mk(), w, h, max_depth=depth)
    else:
        pytest.skip("needs oracle/_ref/libref_rtu.so to build the scene (prebuilt in the build container)")
    check_scene(bufs, karg, cam)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_rtu_golden(path):
    g, karg = load_golden(path)
    check_scene(g, karg, int(g["camera"]), g["hits"].view(R.HIT_DT), g["image"])


def test_rtu_product_builder_vs_oracle():
    """scene prepared by the product's own layout-B host builder (vxh_rtx_*), ~100K triangles"""
    import importlib
    from rt_common import scenes
    from rtu_common import b_mesh
    host = importlib.import_module("vortex-raytracing_b200.host")
    hs = host.RtxHostScene([b_mesh(scenes.heightfield(224), two_materials=True)])
    karg = hs.setup(480, 270)
    check_scene(hs.buffers(), karg, 1)


def test_rtu_full_size_fast_equals_strict():
    """BASELINE configs[1] geometry at 1920x1080 in layout B: the fast order returns the hit records of the reference's
    own visiting order (strict) — a size-independent check that needs no CPU run"""
    import importlib
    from rt_common import scenes
    from rtu_common import b_mesh
    host = importlib.import_module("vortex-raytracing_b200.host")
    hs = host.RtxHostScene([b_mesh(scenes.heightfield(224))])
    c = np.float32(np.cos(np.pi / 4))   # look down on the terrain: RotateX(-pi/4), as the raycast host does (raycast/tracer.cpp:157)
    hs.apply_transform(np.array([[1, 0, 0, 0], [0, c, c, 0], [0, -c, c, 0], [0, 0, 0, 1]], dtype=np.float32))
    karg = hs.setup(1920, 1080)
    tr = tracer_b(hs.buffers(), karg, 1)
    try:
        tr.set_order(1)
        hs_, fs = tr.trace_primary(), tr.run()
        tr.set_order(0)
        hf, ff = tr.trace_primary(), tr.run()
        hp = hit_parity(hf, hs_)
        assert_hit_gates(hp, False)
        assert np.array_equal(hf["dist"], hs_["dist"])
        assert channel_diff(ff, fs) <= 1
        assert (hf["dist"] < 1e30).mean() > 0.2
    finally:
        tr.close()


def test_rtu_any_hit_cut_out():
    """an any-hit shader that answers VX_RT_COMMIT_CONT (program 1: triangles of odd index are cut out) — the path the shipped
    any-hit never takes.  Oracle: the same scene with the odd triangles collapsed to points (Moller-Trumbore rejects them,
    the boxes stay conservative), rendered by the shipped shaders on the GPU and by the port on the CPU."""
    import importlib
    from rt_common import scenes
    from rtu_common import b_mesh
    host = importlib.import_module("vortex-raytracing_b200.host")
    hs = host.RtxHostScene([b_mesh(scenes.heightfield(24), two_materials=True),
                            b_mesh(scenes.icosphere((0, 0.5, 0), 0.4, 2), diffuse=(0.2, 0.6, 0.9), reflectivity=0.5)])
    karg = hs.setup(160, 120, max_depth=3)
    bufs = hs.buffers()
    cut = dict(bufs)
    tri = bufs["tri"].view(np.float32).reshape(-1, 9).copy()
    tri[1::2, 3:6] = tri[1::2, 0:3]
    tri[1::2, 6:9] = tri[1::2, 0:3]
    cut["tri"] = tri.reshape(-1).view(np.uint8)
    tr = tracer_b(bufs, karg, 1)
    tr2 = tracer_b(cut, karg, 1)
    try:
        tr.set_program(1)
        got = tr.run().copy()
        tr2.set_program(0)
        want = tr2.run().copy()
        assert np.array_equal(got, want)
        tr.set_program(0)
        assert not np.array_equal(tr.run(), got)          # the cut-out changes the picture
        port = R.RtuPort()
        port.set_fix_tlas_base(1)
        try:
            frame_port = port.render(R.Image(cut, karg), 1, threads=os.cpu_count() or 1)
        finally:
            port.set_fix_tlas_base(0)
        assert channel_diff(got, frame_port) <= 1
        # strict order cannot ignore a candidate (the reference would re-find it forever): reported as a failed run
        tr.set_program(1)
        tr.set_order(1)
        with pytest.raises(Exception):
            tr.run()
    finally:
        tr.close()
        tr2.close()


def test_rtu_obj_asset_end_to_end(tmp_path):
    """asset path -> scene builder -> kernels: the reference's shipped teapot.obj + a PNG texture loaded by the product's OBJ /
    PNG readers, rendered on the GPU and checked against the oracle; the frame written as PPM reads back."""
    import importlib
    assets = os.path.join(ROOT, "oracle", "_ref", "assets")
    if not os.path.exists(os.path.join(assets, "teapot.obj")):
        pytest.skip("shipped assets not next to oracle/_ref (make -C oracle ref in the build container)")
    host = importlib.import_module("vortex-raytracing_b200.host")
    m = host.load_obj(os.path.join(assets, "teapot.obj"))
    m["materials"] = [{"tex": host.load_png(os.path.join(assets, "ceramic.png"))}]     # tracer.cpp:13 pairs the mesh with ceramic.png
    assert m["tris"].shape[0] > 1000
    hs = host.RtxHostScene([m])
    karg = hs.setup(160, 120)
    bufs = hs.buffers()
    check_scene(bufs, karg, 1)
    tr = tracer_b(bufs, karg, 1)
    try:
        frame = tr.run()
    finally:
        tr.close()
    assert len(np.unique(frame)) > 16                       # a textured, shaded teapot, not a flat colour
    host.write_ppm(str(tmp_path / "teapot.ppm"), frame, binary=True)
    raw = (tmp_path / "teapot.ppm").read_bytes()
    head = b"P6\n160 120\n255\n"
    px = np.frombuffer(raw[len(head):], np.uint8).reshape(120, 160, 3).astype(np.uint32)
    assert np.array_equal((px[..., 0] << 16) + (px[..., 1] << 8) + px[..., 2], frame[::-1])


REF_RTX_APP = os.path.join(ROOT, "oracle", "_ref", "raytracing")


@pytest.mark.skipif(not (os.path.exists(REF_RTX_APP) and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "assets", "teapot.obj"))),
                    reason="the compiled reference raytracing app (oracle/_ref/raytracing + assets) did not travel with the snapshot")
@pytest.mark.parametrize("order", ["1", "0"], ids=["strict", "fast"])
def test_unmodified_reference_raytracing_app_runs_on_b200(tmp_path, order):
    """drop-in proof for the RT-unit host: the reference's own raytracing binary (its Tracer, Scene, BVH4 builder + quantizer,
    tinyobj / stb loaders, DCR writes, SBT and shader uploads; linked against the reference stub libvortex.so) with
    VORTEX_DRIVER=b200.  Its CPU path is commented out in the reference (tracer.cpp:300-315), so the PPM it writes is checked
    against the oracle: the same OBJ through the reference's scene builder (oracle/_ref/libref_rtu.so) rendered by the port."""
    import importlib
    import shutil
    import subprocess
    if not R.have_ref():
        pytest.skip("needs oracle/_ref/libref_rtu.so")
    from rt_common import vx
    host = importlib.import_module("vortex-raytracing_b200.host")
    src = os.path.join(ROOT, "oracle", "_ref", "assets")
    assets = tmp_path / "assets"
    assets.mkdir()
    shutil.copy(os.path.join(src, "ceramic.png"), assets / "ceramic.png")
    (assets / "pot.mtl").write_text("newmtl glaze\nKd 0.8 0.8 0.8\nmap_Kd ceramic.png\n")
    (assets / "pot.obj").write_text("mtllib pot.mtl\nusemtl glaze\n" + open(os.path.join(src, "teapot.obj")).read())
    for name, vma in (("kernel.vxbin", 0x80000000), ("miss.vxbin", 0x80100000), ("closest.vxbin", 0x80200000), ("anyhit.vxbin", 0x80300000)):
        (tmp_path / name).write_bytes(vx.dummy_image(vma))
    w, h = 96, 72
    env = dict(os.environ, VORTEX_DRIVER="b200", VORTEX_B200_ORDER=order, VORTEX_B200_CAMERA="1",
               LD_LIBRARY_PATH=os.path.dirname(vx.LIB_PATH) + ":" + os.path.dirname(REF_RTX_APP) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([REF_RTX_APP, "-m", "pot.obj", "-w", str(w), "-h", str(h), "-o", str(tmp_path / "out.ppm")], cwd=tmp_path, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]
    tok = (tmp_path / "out.ppm").read_text().split()
    assert tok[:4] == ["P3", str(w), str(h), "255"]
    px = np.array(tok[4:], dtype=np.uint32).reshape(h, w, 3)[::-1]            # write_ppm stores the rows bottom-up
    got = (px[..., 0] << 16) + (px[..., 1] << 8) + px[..., 2]
    mesh = host.load_obj(str(assets / "pot.obj"))
    bufs, karg = R.RtuRef().build_scene([mesh], w, h)
    port = R.RtuPort()
    want = port.render(R.Image(bufs, karg), 1, threads=os.cpu_count() or 1)
    background = int(want[0, 0])
    assert (want != background).mean() > 0.1                                   # the teapot is in the picture
    assert channel_diff(got, want) <= 1
    assert (got == want).mean() > 0.999


def test_rtu_schedule_feedback_keeps_the_frame():
    """layout B: the frame is the same in every phase of the schedule feedback (row-major trial frames, descending-cost order,
    the kept order), for the fused kernel and for the shader-program kernel, and equals a run with the feedback disabled"""
    import importlib
    from rt_common import scenes
    from rtu_common import b_mesh
    host = importlib.import_module("vortex-raytracing_b200.host")
    hs = host.RtxHostScene([b_mesh(scenes.heightfield(96), two_materials=True),
                            b_mesh(scenes.icosphere((0, 0.5, 0), 0.4, 2), diffuse=(0.2, 0.6, 0.9), reflectivity=0.5)])
    karg = hs.setup(960, 540, max_depth=2)
    bufs = hs.buffers()
    os.environ["VORTEX_B200_SCHEDULE"] = "0"
    try:
        tr = tracer_b(bufs, karg, 1)
        try:
            want = tr.run().copy()
        finally:
            tr.close()
    finally:
        os.environ.pop("VORTEX_B200_SCHEDULE", None)
    tr = tracer_b(bufs, karg, 1)
    try:
        for program in (None, 0):
            tr.set_program(program)
            for frame in range(6):
                assert np.array_equal(tr.run(), want), (program, frame)
    finally:
        tr.close()


@pytest.mark.parametrize("name", ["cornell64", "multi3_mirror", "hf32"])
def test_rtu_extended_modes_vs_port(name):
    """shadow rays / diffuse bounces on the RT-unit layout (VX_DCR_B200_RT_MODE; no reference counterpart): the CUDA kernels
    against the CPU port's trace_ext_b, whose only intersection primitive is the restated RT-unit traversal.  Strict order
    (the reference's visiting order) must give the port's frame bit for bit, fast order within 1/255."""
    mk, w, h, depth, cam = SCENES_B[name]
    if not R.have_ref():
        pytest.skip("needs oracle/_ref/libref_rtu.so to build the scene")
    bufs, karg = R.RtuRef().build_scene(mk(), w, h, max_depth=max(depth, 3))
    port = R.RtuPort()
    img = R.Image(bufs, karg)
    tr = tracer_b(bufs, karg, cam)
    try:
        base = tr.run().copy()
        for mode in (1, 2, 3):
            want = port.render(img, cam, threads=os.cpu_count() or 1, mode=mode)
            port.set_fix_tlas_base(1)
            try:
                want_fix = port.render(img, cam, threads=os.cpu_count() or 1, mode=mode)
            finally:
                port.set_fix_tlas_base(0)
            tr.set_mode(mode)
            for strict in (1, 0):
                tr.set_order(strict)
                got = tr.run()
                ref = want if strict else want_fix
                d = channel_diff(got, ref)
                assert d <= (0 if strict else 1), (mode, strict, d, int((got != ref).sum()))
                if strict:
                    assert np.array_equal(got, want)
            assert (want != base).any() or name == "one_pixel"      # the mode changes the picture
            tr.set_program(0)
            try:
                with pytest.raises(Exception):
                    tr.run()                                        # shader programs have no extended modes
            finally:
                tr.set_program(None)
        tr.set_mode(0)
        tr.set_order(0)
        assert np.array_equal(tr.run(), base)
    finally:
        tr.close()
