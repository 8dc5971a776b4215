"""python tools/quick_frame.py <workload | rtu_1080p_primary_100k> [launches]: sets the workload up like bench.py and launches its
dominant kernel a few times (the command ncu wraps; VORTEX_B200_SCHEDULE=0 keeps every launch on the same kernel build)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

name = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 6
if name == "rtu_1080p_primary_100k":
    host, rtx, scenes = bench.pkg("host"), bench.pkg("rtx_tracer"), bench.pkg("scenes")
    mesh = scenes.heightfield(224)
    bmesh = {"tris": mesh["tris"], "triex": mesh["triex"], "texid": np.zeros(mesh["tris"].shape[0], np.uint32),
             "materials": [{"tex": mesh["tex"]}], "reflectivity": 0.0}
    hs = host.RtxHostScene([bmesh])
    ax, ay = np.float32(-np.pi / 4), np.float32(np.pi / 4)
    rx = np.array([[1, 0, 0, 0], [0, np.cos(ax), -np.sin(ax), 0], [0, np.sin(ax), np.cos(ax), 0], [0, 0, 0, 1]], dtype=np.float32)
    ry = np.array([[np.cos(ay), 0, np.sin(ay), 0], [0, 1, 0, 0], [-np.sin(ay), 0, np.cos(ay), 0], [0, 0, 0, 1]], dtype=np.float32)
    hs.apply_transform(rx @ ry)
    karg = hs.setup(1920, 1080, spp=1, max_depth=1)
    tr = rtx.RtxTracer(1920, 1080, 1, 1, cuda_device=0).init(hs.buffers())
    tr.setup({k: karg[k][0] for k in karg.dtype.names}, camera=rtx.CAMERA_PINHOLE)
    tr._upload_args()
    for _ in range(n):
        tr.launch()
    print(name, tr.dev.stats()["rays"], tr.dev.last_kernel_ms())
    tr.close()
    sys.exit(0)
wl = bench.WORKLOADS[name]
hs, _ = bench.host_scene(wl)
karg = hs.setup(wl["w"], wl["h"], spp=1, max_depth=wl["depth"])
tr = bench.pkg("tracer").Tracer(wl["w"], wl["h"], 1, wl["depth"], cuda_device=0).init(hs.buffers())
tr.setup({k: karg[k][0] for k in karg.dtype.names})
tr.set_mode(wl["mode"], 0)
tr._upload_args()
if "rays_per_gpu" in wl:
    R = int(wl["rays_per_gpu"])
    s = np.float32(wl.get("scale", 1.0))
    rays = bench.pkg("scenes").random_rays(R, tuple(s * np.float32(v) for v in (-1.0, -0.05, -1.0)), tuple(s * np.float32(v) for v in (1.0, 0.45, 1.0)), seed=1)
    rb, hb = tr.dev.upload_bytes(rays), tr.dev.mem_alloc(R * 24)
    for _ in range(n):
        tr.dev.trace_rays(tr.args, rb, R, hb)
        tr.dev.ready_wait()
else:
    for _ in range(n):
        tr.launch()
print(name, tr.dev.stats()["rays"], tr.dev.last_kernel_ms())
tr.close()
