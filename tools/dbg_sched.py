import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
host, rtx, scenes = bench.pkg("host"), bench.pkg("rtx_tracer"), bench.pkg("scenes")
mesh = scenes.heightfield(224)
bm = {"tris": mesh["tris"], "triex": mesh["triex"], "texid": np.zeros(mesh["tris"].shape[0], np.uint32), "materials": [{"tex": mesh["tex"]}], "reflectivity": 0.0}
ax, ay = np.float32(-np.pi / 4), np.float32(np.pi / 4)
rx = np.array([[1, 0, 0, 0], [0, np.cos(ax), -np.sin(ax), 0], [0, np.sin(ax), np.cos(ax), 0], [0, 0, 0, 1]], dtype=np.float32)
ry = np.array([[np.cos(ay), 0, np.sin(ay), 0], [0, 1, 0, 0], [-np.sin(ay), 0, np.cos(ay), 0], [0, 0, 0, 1]], dtype=np.float32)
w, h = 1920, 1080
hs = host.RtxHostScene([bm]); hs.apply_transform(rx @ ry); karg = hs.setup(w, h)
tr = rtx.RtxTracer(w, h, 1, 1).init(hs.buffers())
tr.setup({k: karg[k][0] for k in karg.dtype.names}, camera=rtx.CAMERA_PINHOLE)
tr._upload_args()
for i in range(5):
    tr.launch(); print("frame", i, tr.dev.last_kernel_ms(), flush=True)
tr.close()
