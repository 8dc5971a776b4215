"""host-side breakdown of one end-to-end frame of cfg1 (the bench's e2e step): seconds per vx_* call, averaged"""
import os
import sys
import time

import numpy as np

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import bench  # noqa: E402
import torch  # noqa: E402

wl = bench.WORKLOADS["cfg1_1080p_primary+shadow_100k"]
tracer_mod, host = bench.pkg("tracer"), bench.pkg("host")
hs = host.HostScene(bench.make_meshes(wl))
W, H = wl["w"], wl["h"]
karg = hs.setup(W, H, spp=1, max_depth=wl["depth"])
bufs = hs.buffers()
tr = tracer_mod.Tracer(W, H, 1, wl["depth"], cuda_device=0).init(bufs)
tr.setup({k: karg[k][0] for k in karg.dtype.names})
tr.set_mode(wl["mode"], 0)
pinned = {n: torch.from_numpy(bufs[n].copy()).pin_memory() for n in tracer_mod.SCENE_BUFFERS if bufs[n].nbytes}
out_pinned = torch.empty((H, W), dtype=torch.int32).pin_memory()
acc = {}


def timed(name, fn):
    t0 = time.perf_counter()
    fn()
    acc[name] = acc.get(name, 0.0) + time.perf_counter() - t0


N = 30
for it in range(N + 5):
    if it == 5:
        acc.clear()
        t_all = time.perf_counter()
    for n, t in pinned.items():
        timed(f"copy_to_dev[{n}] {t.numel() / 1e6:.2f} MB", lambda: tr.bufs[n].copy_to_dev((t.data_ptr(), t.numel())))
    timed("upload_args", tr._upload_args)
    timed("vx_start", lambda: tr.dev.start(tr.krnl, tr.args))
    timed("vx_ready_wait", tr.dev.ready_wait)
    timed("copy_from_dev 8.29 MB", lambda: tr.out.copy_from_dev((out_pinned.data_ptr(), W * H * 4)))
total = time.perf_counter() - t_all
for k, v in acc.items():
    print(f"{k:34s} {v / N * 1e6:9.1f} us")
print(f"{'sum of calls':34s} {sum(acc.values()) / N * 1e6:9.1f} us   wall per frame {total / N * 1e6:.1f} us   kernel {tr.dev.last_kernel_ms() * 1e3:.1f} us")
st = tr.dev.stats()
print(f"last run: prep_ms {st['prep_ms']*1e3:.1f} us ({st['prep_launches']} launches), kernel_ms {st['kernel_ms']*1e3:.1f} us")
tr.close()
