"""Shared helpers for the tests: package import, oracle selection, parity metrics."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pkg = importlib.import_module("vortex-raytracing_b200")
scenes = importlib.import_module("vortex-raytracing_b200.scenes")
tracer_mod = importlib.import_module("vortex-raytracing_b200.tracer")
vx = importlib.import_module("vortex-raytracing_b200.vx")


def pyoracle():
    return importlib.import_module("oracle.pyoracle")


def best_oracle():
    return pyoracle().best()


parity = importlib.import_module("vortex-raytracing_b200.parity")
channel_diff, hit_parity = parity.channel_diff, parity.hit_parity


def tracer_for(sc, karg, cuda_device=None, gpus=None):
    """A vx_* Tracer holding exactly the bytes of oracle scene `sc` (after sc.setup -> karg)."""
    w, h = int(karg["dst_width"][0]), int(karg["dst_height"][0])
    tr = tracer_mod.Tracer(w, h, int(karg["samples_per_pixel"][0]), int(karg["max_depth"][0]), cuda_device=cuda_device, gpus=gpus)
    bufs = sc.buffers() if hasattr(sc, "buffers") else sc.bufs
    tr.init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    return tr


def assert_hit_gates(hp, strict):
    """north_star gates: IDs bit-exact on >= 99.99 % of rays, residue = equal-t ties; t within 1e-5 relative."""
    assert hp["non_tie"] == 0, hp
    assert hp["mismatch_frac"] <= 1e-4, hp
    assert hp["max_rel_t"] <= 1e-5, hp
    if strict:
        assert hp["mismatch"] == 0 and hp["bit_exact_t"] == hp["n"], hp
