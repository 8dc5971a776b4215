"""follow-up 2: in the end-to-end frame the 6.4 MB bvh upload takes 212 us where 128 us is the link rate; which of the
preceding calls causes it?"""
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


def up(n):
    t = pinned[n]
    t0 = time.perf_counter()
    tr.bufs[n].copy_to_dev((t.data_ptr(), t.numel() * t.element_size()))
    return (time.perf_counter() - t0) * 1e6


def frame(download):
    tr._upload_args()
    tr.dev.start(tr.krnl, tr.args)
    tr.dev.ready_wait()
    if download:
        tr.out.copy_from_dev((out_pinned.data_ptr(), W * H * 4))


for download in (False, True):
    for order in (("tri", "triEx", "triIdx", "tlas", "blas", "bvh", "tex"), ("tri", "triEx", "bvh"), ("tri", "triEx", "triIdx", "bvh"),
                  ("tri", "triEx", "tlas", "bvh"), ("bvh", "tri", "triEx", "triIdx", "tlas", "blas", "tex"), ("tri", "bvh", "triEx")):
        acc = {n: [] for n in order}
        for i in range(25):
            for n in order:
                acc[n].append(up(n))
            frame(download)
        print(f"download={download}: " + "  ".join(f"{n} {np.median(acc[n][5:]):.0f}" for n in order))
tr.close()
