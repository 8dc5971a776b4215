"""Where the end-to-end frame time goes on an N-GPU device (python tools/r02_e2e_probe.py N):
 (1) raw link probe: concurrent pinned-host -> device / device -> pinned-host copies of one slice per GPU (torch streams);
 (2) the vx call sequence of bench.py's e2e leg, each phase timed on its own: uploads only, start+wait only, download only."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1
name = sys.argv[2] if len(sys.argv) > 2 else bench.DEFAULT_WORKLOAD


def raw_probe(total_mb=160):
    per = (total_mb << 20) // N
    host = torch.empty(per * N, dtype=torch.uint8).pin_memory()
    devs = [torch.empty(per, dtype=torch.uint8, device=f"cuda:{i}") for i in range(N)]
    streams = [torch.cuda.Stream(device=i) for i in range(N)]

    def run(h2d):
        for i in range(N):
            with torch.cuda.stream(streams[i]):
                if h2d:
                    devs[i].copy_(host[i * per:(i + 1) * per], non_blocking=True)
                else:
                    host[i * per:(i + 1) * per].copy_(devs[i], non_blocking=True)
        for s in streams:
            s.synchronize()
    out = {}
    for h2d in (True, False):
        for _ in range(5):
            run(h2d)
        t0 = time.perf_counter()
        for _ in range(10):
            run(h2d)
        dt = (time.perf_counter() - t0) / 10
        out["h2d" if h2d else "d2h"] = (per * N / dt / 1e9, dt * 1e3)
    return out


print(f"N={N} raw link probe (one slice per GPU, {160 // N} MB each, all concurrently): " +
      ", ".join(f"{k} {v[0]:.1f} GB/s ({v[1]:.3f} ms)" for k, v in raw_probe().items()), flush=True)
print(f"N={N} raw link probe, 36 MB in all (the size of the tri buffer; {36 / N:.1f} MB per GPU): " +
      ", ".join(f"{k} {v[0]:.1f} GB/s ({v[1]:.3f} ms)" for k, v in raw_probe(36).items()), flush=True)

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
for _ in range(5):
    dev.start(tr.krnl, tr.args)
    dev.ready_wait()


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


def uploads():
    for n, t in pinned.items():
        tr.bufs[n].copy_to_dev((t.data_ptr(), t.numel()))


def uploads_timed():
    out = {}
    sync()
    for n, t in pinned.items():
        t0 = time.perf_counter()
        tr.bufs[n].copy_to_dev((t.data_ptr(), t.numel()))
        out[n] = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    sync()
    out["drain"] = (time.perf_counter() - t0) * 1e3
    return out


def run():
    dev.start(tr.krnl, tr.args)
    dev.ready_wait()


def download():
    tr.out.copy_from_dev((out_pinned.data_ptr(), W * H * 4))


def full():
    uploads()
    tr._upload_args()
    run()
    download()


mb_up = sum(t.numel() for t in pinned.values()) / 1e6
print(f"uploads ({mb_up:.1f} MB): {phase(uploads):.3f} ms; per call (ms): " + ", ".join(f"{k} {v:.3f}" for k, v in uploads_timed().items()))
print(f"start+wait: {phase(run):.3f} ms   (kernel_ms {dev.stats()['kernel_ms']:.3f}, group_ms {dev.stats()['group_ms']:.3f})")
run()
print(f"download ({W * H * 4 / 1e6:.1f} MB): {phase(download):.3f} ms")
print(f"full sequence: {phase(full):.3f} ms")
dev.set_readback(tr.out, out_pinned.data_ptr(), W * H * 4)
print(f"start+wait with readback (no uploads): {phase(run):.3f} ms   (kernel_ms {dev.stats()['kernel_ms']:.3f})")
print(f"full sequence with readback: {phase(full):.3f} ms   (launches per step {dev.stats()['kernel_launches']}, kernel_ms {dev.stats()['kernel_ms']:.3f})")
tr.close()
