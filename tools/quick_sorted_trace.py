"""does ordering the incoherent rays help? kernel time of vx_b200_trace_rays on cfg4 rays as generated vs sorted on the host by
(direction octant, Morton code of the origin) — an experiment before building a device-side sort.  python tools/quick_sorted_trace.py"""
import os
import sys

import numpy as np

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import bench  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg4_incoherent_10m"
wl = bench.WORKLOADS[name]
tracer_mod, host, scenes = bench.pkg("tracer"), bench.pkg("host"), bench.pkg("scenes")
R = int(wl["rays_per_gpu"])
hs = host.HostScene(bench.make_meshes(wl))
karg = hs.setup(wl["w"], wl["h"], spp=1, max_depth=1)
tr = tracer_mod.Tracer(wl["w"], wl["h"], 1, 1, cuda_device=0).init(hs.buffers())
tr.setup({k: karg[k][0] for k in karg.dtype.names})
tr.set_mode(0, 0)
tr._upload_args()
s = np.float32(wl.get("scale", 1.0))
lo = np.array([s * np.float32(v) for v in (-1.0, -0.05, -1.0)], np.float32)
hi = np.array([s * np.float32(v) for v in (1.0, 0.45, 1.0)], np.float32)
rays = scenes.random_rays(R, tuple(lo), tuple(hi), seed=1)


def spread(v):
    v = v.astype(np.uint64) & 0x3FF
    v = (v | (v << 16)) & 0x30000FF
    v = (v | (v << 8)) & 0x300F00F
    v = (v | (v << 4)) & 0x30C30C3
    v = (v | (v << 2)) & 0x9249249
    return v


def keys(r, bits_dir):
    q = np.clip(((r[:, :3] - lo) / (hi - lo) * 1023.0), 0, 1023).astype(np.uint32)
    m = (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2])
    octant = ((r[:, 3] < 0).astype(np.uint64) << 2) | ((r[:, 4] < 0).astype(np.uint64) << 1) | (r[:, 5] < 0).astype(np.uint64)
    if bits_dir == "octant_high":
        return (octant << 30) | m
    if bits_dir == "octant_low":       # cells of 2^-5 per axis first, octant inside the cell, then the finer Morton bits
        return ((m >> 15) << 18) | (octant << 15) | (m & 0x7FFF)
    return m


hb = tr.dev.mem_alloc(R * 24)
for label in ("as generated", "origin", "octant_high", "octant_low"):
    r = rays if label == "as generated" else rays[np.argsort(keys(rays, label), kind="stable")]
    rb = tr.dev.upload_bytes(np.ascontiguousarray(r))
    ms = []
    for i in range(6):
        tr.dev.trace_rays(tr.args, rb, R, hb)
        tr.dev.ready_wait()
        ms.append(tr.dev.last_kernel_ms())
    tr.dev.enable_stats(True)
    tr.dev.trace_rays(tr.args, rb, R, hb)
    tr.dev.ready_wait()
    st = tr.dev.stats()
    tr.dev.enable_stats(False)
    print(f"{label:14s} kernel_ms={float(np.mean(ms[2:])):.3f} Mrays/s={R / float(np.mean(ms[2:])) / 1e3:.0f} simt={st['lane_steps'] / (32.0 * st['warp_steps']):.3f}", flush=True)
    rb.free()
tr.close()
