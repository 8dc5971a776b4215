"""cfg4 (incoherent rays, 10 M triangles): scheduling parameters of the hit-record state machine with the stack ring
(python tools/r02_sweep_trace_quorum.py)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

wl = bench.WORKLOADS["cfg4_incoherent_10m"]
hs, _ = bench.host_scene(wl)
karg = hs.setup(8, 8, spp=1, max_depth=1)
tr = bench.pkg("tracer").Tracer(8, 8, 1, 1, cuda_device=0).init(hs.buffers())
tr.setup({k: karg[k][0] for k in karg.dtype.names})
tr.set_mode(0, 0)
tr._upload_args()
R = 1 << 23
s = np.float32(wl["scale"])
rays = bench.pkg("scenes").random_rays(R, tuple(s * np.float32(v) for v in (-1.0, -0.05, -1.0)), tuple(s * np.float32(v) for v in (1.0, 0.45, 1.0)), seed=1)
rb, hb = tr.dev.upload_bytes(rays), tr.dev.mem_alloc(R * 24)


def run():
    ms = []
    for i in range(9):
        tr.dev.flush_l2()
        tr.dev.trace_rays(tr.args, rb, R, hb)
        tr.dev.ready_wait()
        if i >= 3:
            ms.append(tr.dev.last_kernel_ms())
    return R / float(np.mean(ms)) / 1e3


for refill, inner, chunk in ((16, 12, 64), (12, 8, 64), (16, 8, 64), (20, 12, 64), (20, 16, 64), (24, 16, 64), (24, 20, 64), (28, 24, 64), (16, 12, 32), (16, 12, 128),
                             (16, 12, 64)):
    os.environ["VORTEX_B200_TRACE_QUORUM_REFILL"], os.environ["VORTEX_B200_TRACE_QUORUM_INNER"], os.environ["VORTEX_B200_TRACE_CHUNK"] = str(refill), str(inner), str(chunk)
    print(f"refill {refill} inner {inner} chunk {chunk}: {run():.0f} Mrays/s", flush=True)
tr.close()
