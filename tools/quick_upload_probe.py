"""why does one scene buffer upload slower than the others?  times copy_to_dev of every cfg1 scene buffer alone, back to back,
and right after a frame was rendered.  python tools/quick_upload_probe.py"""
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
pinned = {n: torch.from_numpy(bufs[n].copy()).pin_memory() for n in ("tri", "triEx", "bvh")}
big = torch.empty(8 << 20, dtype=torch.uint8).pin_memory()


def up(n, t=None):
    t = pinned[n] if t is None else t
    nb = pinned[n].numel() * pinned[n].element_size()
    t0 = time.perf_counter()
    tr.bufs[n].copy_to_dev((t.data_ptr(), nb))
    return (time.perf_counter() - t0) * 1e6, nb


def frame():
    tr._upload_args()
    tr.dev.start(tr.krnl, tr.args)
    tr.dev.ready_wait()


for label, pre in (("back to back", None), ("after a rendered frame", frame)):
    print(label)
    for n in ("tri", "triEx", "bvh", "bvh", "triEx", "tri"):
        v = []
        for i in range(12):
            if pre:
                pre()
            us, nb = up(n)
            v.append(us)
        print(f"  {n:6s} {nb / 1e6:6.2f} MB  median {np.median(v[2:]):7.1f} us  {nb / np.median(v[2:]) / 1e3:5.1f} GB/s  min {min(v):7.1f}")
print("same destinations, one shared 8 MB pinned source")
for n in ("tri", "triEx", "bvh"):
    v = [up(n, big)[0] for i in range(12)]
    print(f"  {n:6s} median {np.median(v[2:]):7.1f} us")
print("frame, then all three in the bench's order (tri, triEx, bvh) and in reverse")
for order in (("tri", "triEx", "bvh"), ("bvh", "triEx", "tri")):
    acc = {n: [] for n in order}
    for i in range(12):
        frame()
        for n in order:
            acc[n].append(up(n)[0])
    print("  " + "  ".join(f"{n} {np.median(acc[n][2:]):.1f} us" for n in order))
tr.close()
