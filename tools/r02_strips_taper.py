"""Strip count x taper of the strip-pipelined readback on one workload, one process (the library reads both knobs at every
vx_start): python tools/r02_strips_taper.py [N] [workload].  Prints start+wait and the full upload/start/wait/download sequence."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1
name = sys.argv[2] if len(sys.argv) > 2 else bench.DEFAULT_WORKLOAD
tracer_mod = bench.pkg("tracer")
wl = bench.WORKLOADS[name]
W, H = wl["w"], wl["h"]
hs, _ = bench.host_scene(wl)
karg = hs.setup(W, H, spp=1, max_depth=wl["depth"])
bufs = hs.buffers()
tr = tracer_mod.Tracer(W, H, 1, wl["depth"], cuda_device=0, gpus=N).init(bufs)
tr.setup({k: karg[k][0] for k in karg.dtype.names})
tr.set_mode(wl["mode"], 0)
tr._upload_args()
dev = tr.dev
pinned = {n: torch.from_numpy(bufs[n].copy()).pin_memory() for n in tracer_mod.SCENE_BUFFERS if bufs[n].nbytes}
out_pinned = torch.empty((H, W), dtype=torch.int32).pin_memory()


def sync():
    for i in range(N):
        torch.cuda.synchronize(i)


def phase(fn, n=10):
    for _ in range(3):
        fn()
    sync()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    sync()
    return (time.perf_counter() - t0) / n * 1e3


def run():
    dev.start(tr.krnl, tr.args)
    dev.ready_wait()


def run_dl():
    run()
    tr.out.copy_from_dev((out_pinned.data_ptr(), W * H * 4))


def full():
    for n, t in pinned.items():
        tr.bufs[n].copy_to_dev((t.data_ptr(), t.numel()))
    tr._upload_args()
    run_dl()


print(f"{name} N={N}: plain start+wait {phase(run):.3f} ms, + download {phase(run_dl):.3f} ms, full sequence {phase(full):.3f} ms", flush=True)
dev.set_readback(tr.out, out_pinned.data_ptr(), W * H * 4)
ref = None
VARIANTS = [(16, 90, 0), (16, 90, 2), (16, 90, 1), (16, 90, 2), (8, 100, 0), (8, 100, 2), (16, 80, 2), (12, 90, 2)]   # (strips, taper %, order: 0 top-down, 1 bottom-up, 2 by feedback)
for strips, taper, order in VARIANTS:
    os.environ["VORTEX_B200_READBACK_STRIPS"] = str(strips)
    os.environ["VORTEX_B200_READBACK_TAPER"] = str(taper)
    os.environ["VORTEX_B200_READBACK_ORDER"] = str(order)
    a, b = phase(run_dl), phase(full)
    out_pinned.zero_()   # the comparison below must see THIS variant's copy
    run_dl()
    sync()
    img = out_pinned.clone()
    same = "first" if ref is None else ("same image" if bool((img == ref).all()) else "IMAGE DIFFERS")
    ref = img if ref is None else ref
    print(f"strips {strips:2d} taper {taper:3d} order {order}: start+wait+download {a:.3f} ms, full sequence {b:.3f} ms  (kernel_ms {dev.stats()['kernel_ms']:.3f}, launches {dev.stats()['kernel_launches']}, {same})", flush=True)
tr.close()
