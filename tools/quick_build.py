"""GPU vs host BVH build time and the tracing cost of the resulting trees (scratch tool): python tools/quick_build.py [n]"""
import importlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
scenes = importlib.import_module("vortex-raytracing_b200.scenes")
host = importlib.import_module("vortex-raytracing_b200.host")
tracer_mod = importlib.import_module("vortex-raytracing_b200.tracer")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 708
w, h = 3840, 2160
mesh = scenes.heightfield(n)
t0 = time.perf_counter()
hs = host.HostScene([mesh])
karg = hs.setup(w, h)
bufs = hs.buffers()
host_s = time.perf_counter() - t0
tr = tracer_mod.Tracer(w, h, 1, 1).init(bufs)
tr.setup({k: karg[k][0] for k in karg.dtype.names})
tr.set_mode(0, 0)
tr._upload_args()


def frame_ms():
    ms = []
    for i in range(6):
        tr.launch()
        if i >= 2:
            ms.append(tr.dev.last_kernel_ms())
    return float(np.mean(ms))


host_ms = frame_ms()
nt = mesh["tris"].shape[0]
print(f"triangles={nt} host_build_s={host_s:.3f} primary_4k_frame_ms host SAH tree={host_ms:.3f}")
for leaf, ctrav in ((1, 1.0), (2, 0.0), (2, 1.0), (4, 0.0), (4, 0.5), (4, 1.0), (4, 2.0), (8, 1.0), (8, 2.0), (14, 2.0)):
    os.environ["VORTEX_B200_BUILD_MAX_LEAF"], os.environ["VORTEX_B200_BUILD_CTRAV"] = str(leaf), str(ctrav)
    tr.dev.build_bvh(tr.bufs["tri"], nt, tr.bufs["bvh"], tr.bufs["triIdx"])
    tr.dev.build_bvh(tr.bufs["tri"], nt, tr.bufs["bvh"], tr.bufs["triIdx"])
    gpu_build_ms = tr.dev.stats()["kernel_ms"]
    gpu_ms = frame_ms()
    print(f"  GPU build max_leaf={leaf} c_trav={ctrav}: build {gpu_build_ms:.2f} ms, frame {gpu_ms:.3f} ms = {gpu_ms / host_ms:.3f} x the SAH tree's")
os.environ.pop("VORTEX_B200_BUILD_MAX_LEAF"), os.environ.pop("VORTEX_B200_BUILD_CTRAV")
tr.close()
if os.environ.get("LAYOUT_A_ONLY"):
    sys.exit(0)

# ---- layout B: quantized BVH4 (vx_b200_build_qbvh) ----
rtx = importlib.import_module("vortex-raytracing_b200.rtx_tracer")
bm = {"tris": mesh["tris"], "triex": mesh["triex"], "texid": np.zeros(nt, np.uint32), "materials": [{"tex": mesh["tex"]}], "reflectivity": 0.0}
t0 = time.perf_counter()
hb = host.RtxHostScene([bm])
c = np.float32(np.cos(np.pi / 4))
hb.apply_transform(np.array([[1, 0, 0, 0], [0, c, c, 0], [0, -c, c, 0], [0, 0, 0, 1]], dtype=np.float32))
kb = hb.setup(w, h)
bb = hb.buffers()
host_b_s = time.perf_counter() - t0
tb = rtx.RtxTracer(w, h, 1, 1).init(bb)
tb.setup({k: kb[k][0] for k in kb.dtype.names}, camera=rtx.CAMERA_PINHOLE)
tb._upload_args()


def frame_b_ms():
    ms = []
    for i in range(6):
        tb.launch()
        if i >= 2:
            ms.append(tb.dev.last_kernel_ms())
    return float(np.mean(ms))


host_b_ms = frame_b_ms()
tb.dev.build_qbvh(tb.bufs["tri"], nt, tb.bufs["triEx"], tb.bufs["qbvh"])
used, _ = tb.dev.build_qbvh(tb.bufs["tri"], nt, tb.bufs["triEx"], tb.bufs["qbvh"])
gpu_b_build_ms = tb.dev.stats()["kernel_ms"]
gpu_b_ms = frame_b_ms()
print(f"layout B: triangles={nt} host_build_s={host_b_s:.3f} (nodes {bb['bvh_nodes_used']}) gpu_build_ms={gpu_b_build_ms:.2f} (nodes {used}) "
      f"primary_4k_frame_ms host_tree={host_b_ms:.3f} gpu_tree={gpu_b_ms:.3f}")
tb.close()
