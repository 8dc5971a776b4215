"""python tools/dump_hits.py <lib.so | -> <workload> <out.npy>: closest-hit records of every primary ray of the workload's frame
(vx_b200_trace_primary, fast order) — for GPU-vs-GPU comparisons of kernel variants over whole frames."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

lib, name, out = sys.argv[1:4]
if lib != "-":
    bench.pkg("vx").LIB_PATH = os.path.abspath(lib)
wl = bench.WORKLOADS[name]
hs, _ = bench.host_scene(wl)
karg = hs.setup(wl["w"], wl["h"], spp=1, max_depth=wl["depth"])
tr = bench.pkg("tracer").Tracer(wl["w"], wl["h"], 1, wl["depth"]).init(hs.buffers())
tr.setup({k: karg[k][0] for k in karg.dtype.names})
tr.set_mode(wl["mode"], 0)
hits = tr.trace_primary()
img = tr.run()
np.save(out, hits)
np.save(out + ".img.npy", img)
tr.close()
print(lib, name, "hits", len(hits), "hit fraction", float((hits["dist"] < 1e30).mean()))
