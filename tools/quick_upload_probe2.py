"""follow-up to quick_upload_probe.py: time evolution of back-to-back uploads, and whether a kernel before each upload changes it"""
import os
import sys
import time

import numpy as np

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import bench  # noqa: E402
import torch  # noqa: E402

tracer_mod, host = bench.pkg("tracer"), bench.pkg("host")
vx = bench.pkg("vx")
dev = vx.Device(0) if hasattr(vx, "Device") else None
NB = 6422528
buf = dev.mem_alloc(NB)
src = torch.empty(NB, dtype=torch.uint8).pin_memory()
src.fill_(3)
scratch = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
one = torch.zeros(32, device="cuda")


def up():
    t0 = time.perf_counter()
    buf.copy_to_dev((src.data_ptr(), NB))
    return (time.perf_counter() - t0) * 1e6


def series(label, pre, n=40):
    v = []
    for i in range(n):
        if pre:
            pre()
        v.append(up())
    print(f"{label:44s} " + " ".join(f"{x:.0f}" for x in v))


def tiny():
    one.add_(1.0)
    torch.cuda.synchronize()


def fill():
    scratch.fill_(1)
    torch.cuda.synchronize()


def tiny_nosync():
    one.add_(1.0)


series("back to back", None)
series("tiny kernel + sync before each", tiny)
series("back to back again", None)
series("64 MB fill + sync before each", fill)
series("tiny kernel, no sync, before each", tiny_nosync)
series("sleep 1 ms before each", lambda: time.sleep(0.001))
series("back to back, last", None)
