"""fixed cost of one trace launch: kernel time of vx_b200_trace_rays for small ray counts (machine and simple kernels)"""
import os
import sys

import numpy as np

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import bench  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg4_incoherent_10m"
wl = bench.WORKLOADS[name]
tracer_mod, host, scenes, vx = bench.pkg("tracer"), bench.pkg("host"), bench.pkg("scenes"), bench.pkg("vx")
R = 1 << 20
hs = host.HostScene(bench.make_meshes(wl))
karg = hs.setup(wl["w"], wl["h"], spp=1, max_depth=1)
tr = tracer_mod.Tracer(wl["w"], wl["h"], 1, 1, cuda_device=0).init(hs.buffers())
tr.setup({k: karg[k][0] for k in karg.dtype.names})
tr._upload_args()
s = np.float32(wl.get("scale", 1.0))
rays = scenes.random_rays(R, tuple(s * np.float32(v) for v in (-1.0, -0.05, -1.0)), tuple(s * np.float32(v) for v in (1.0, 0.45, 1.0)), seed=1)
rb = tr.dev.upload_bytes(rays)
hb = tr.dev.mem_alloc(R * 24)
for kernel in (0, 1):
    tr.set_mode(0, 0, kernel=kernel)
    print("kernel", "machine" if kernel == 0 else "one ray per lane")
    for n in (32, 1024, 8192, 32768, 65536, 131072, 262144, 524288, 1 << 20):
        ms = []
        for i in range(6):
            tr.dev.trace_rays(tr.args, rb, n, hb)
            tr.dev.ready_wait()
            ms.append(tr.dev.last_kernel_ms())
        print(f"  n={n:8d} kernel_ms={float(np.mean(ms[2:])):.4f} min={min(ms[2:]):.4f}")
tr.close()
