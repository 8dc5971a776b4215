"""fixed cost of one layout-B frame launch: kernel time of k_rtu_render on the cfg1 geometry at several resolutions"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

host, rtx, scenes = bench.pkg("host"), bench.pkg("rtx_tracer"), bench.pkg("scenes")
mesh = scenes.heightfield(224)
bm = {"tris": mesh["tris"], "triex": mesh["triex"], "texid": np.zeros(mesh["tris"].shape[0], np.uint32), "materials": [{"tex": mesh["tex"]}], "reflectivity": 0.0}
ax, ay = np.float32(-np.pi / 4), np.float32(np.pi / 4)
rx = np.array([[1, 0, 0, 0], [0, np.cos(ax), -np.sin(ax), 0], [0, np.sin(ax), np.cos(ax), 0], [0, 0, 0, 1]], dtype=np.float32)
ry = np.array([[np.cos(ay), 0, np.sin(ay), 0], [0, 1, 0, 0], [-np.sin(ay), 0, np.cos(ay), 0], [0, 0, 0, 1]], dtype=np.float32)
for w, h in ((8, 8), (240, 135), (480, 270), (960, 540), (1920, 1080), (3840, 2160), (7680, 4320)):
    hs = host.RtxHostScene([bm])
    hs.apply_transform(rx @ ry)
    karg = hs.setup(w, h)
    tr = rtx.RtxTracer(w, h, 1, 1).init(hs.buffers())
    tr.setup({k: karg[k][0] for k in karg.dtype.names}, camera=rtx.CAMERA_PINHOLE)
    tr._upload_args()
    out = []
    for prog in (None, 0):
        tr.set_program(prog)
        ms = []
        for i in range(8):
            tr.launch()
            ms.append(tr.dev.last_kernel_ms())
        out.append(min(ms[3:]))
    print(f"{w}x{h}: fused kernel_ms={out[0]:.4f} ({w * h / out[0] / 1e3:.0f} Mrays/s)  program kernel_ms={out[1]:.4f}", flush=True)
    tr.close()
