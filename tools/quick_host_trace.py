"""where the time of vx_b200_trace_rays_host goes: wall clock, summed kernel time and launches per chunk size, next to the
device-buffer kernel on the same ray counts.  python tools/quick_host_trace.py [workload]"""
import os
import sys
import time

import numpy as np

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import bench  # noqa: E402
import torch  # noqa: E402

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
rays = scenes.random_rays(R, tuple(s * np.float32(v) for v in (-1.0, -0.05, -1.0)), tuple(s * np.float32(v) for v in (1.0, 0.45, 1.0)), seed=1)
rp = torch.from_numpy(rays).pin_memory()
hp = torch.empty(R * 24, dtype=torch.uint8).pin_memory()
rb = tr.dev.mem_alloc(R * 24)
hb = tr.dev.mem_alloc(R * 24)
rb.copy_to_dev((rp.data_ptr(), R * 24))
print("device-buffer kernel by ray count")
for n in (1 << 17, 1 << 18, 1 << 19, 1 << 20, 1 << 21, 1 << 22, R):
    ms = []
    for i in range(5):
        tr.dev.trace_rays(tr.args, rb, n, hb)
        tr.dev.ready_wait()
        ms.append(tr.dev.last_kernel_ms())
    k = float(np.mean(ms[2:]))
    print(f"  n={n:9d} kernel_ms={k:.4f} Mrays/s={n / k / 1e3:.0f}")
print("copies alone (pinned)")
for n in (1 << 18, 1 << 20, R):
    for _ in range(2):
        t0 = time.perf_counter()
        rb.copy_to_dev((rp.data_ptr(), n * 24))
        t1 = time.perf_counter()
        hb.copy_from_dev((hp.data_ptr(), n * 24))
        t2 = time.perf_counter()
    print(f"  n={n:9d} h2d_ms={(t1 - t0) * 1e3:.3f} ({n * 24 / (t1 - t0) / 1e9:.1f} GB/s) d2h_ms={(t2 - t1) * 1e3:.3f} ({n * 24 / (t2 - t1) / 1e9:.1f} GB/s)")
print("vx_b200_trace_rays_host by chunk and resident blocks per SM of each chunk's kernel")
for bps in (1, 2, 4):
    os.environ["VORTEX_B200_HOST_BLOCKS_PER_SM"] = str(bps)
    for chunk in (1 << 17, 1 << 18, 1 << 19, 1 << 20, 1 << 21):
        os.environ["VORTEX_B200_HOST_CHUNK"] = str(chunk)
        best = None
        for i in range(4):
            t0 = time.perf_counter()
            tr.dev.trace_rays_host(tr.args, rp.data_ptr(), R, hp.data_ptr())
            dt = time.perf_counter() - t0
            if i and (best is None or dt < best):
                best = dt
        st = tr.dev.stats()
        print(f"  blocks/SM={bps} chunk={chunk:8d} wall_ms={best * 1e3:.3f} Mrays/s={R / best / 1e6:.0f} sum_kernel_ms={st['kernel_ms']:.3f} launches={st['kernel_launches']}")
tr.close()
