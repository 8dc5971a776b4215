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


def channel_diff(a, b):
    """max per-channel |a-b| of two 0x00RRGGBB images"""
    d = 0
    for sh in (16, 8, 0):
        ca = ((a >> sh) & 255).astype(np.int32)
        cb = ((b >> sh) & 255).astype(np.int32)
        d = max(d, int(np.abs(ca - cb).max()))
    return d


def hit_parity(gpu, ref):
    """ID mismatch statistics between two ray_hit_t arrays (north_star gates)."""
    n = len(ref)
    miss_g, miss_r = gpu["dist"] == np.float32(1e30), ref["dist"] == np.float32(1e30)
    same = (miss_g & miss_r) | ((~miss_g) & (~miss_r) & (gpu["triIdx"] == ref["triIdx"]) & (gpu["blasIdx"] == ref["blasIdx"]))
    bad = np.flatnonzero(~same)
    both = (~miss_g) & (~miss_r)
    rel = np.zeros(n, dtype=np.float64)
    rel[both] = np.abs(gpu["dist"][both].astype(np.float64) - ref["dist"][both]) / np.maximum(np.abs(ref["dist"][both]), 1e-30)
    # a mismatch is a "tie" when both sides hit and report (nearly) the same distance
    tie = np.array([both[i] and rel[i] <= 1e-5 for i in bad], dtype=bool) if len(bad) else np.zeros(0, bool)
    return {"n": n, "mismatch": len(bad), "mismatch_frac": len(bad) / max(n, 1), "ties": int(tie.sum()),
            "non_tie": int(len(bad) - tie.sum()), "max_rel_t": float(rel[same & both].max()) if (same & both).any() else 0.0,
            "bit_exact_t": int((gpu["dist"][same] == ref["dist"][same]).sum()), "same": int(same.sum())}


def tracer_for(sc, karg, cuda_device=None):
    """A vx_* Tracer holding exactly the bytes of oracle scene `sc` (after sc.setup -> karg)."""
    w, h = int(karg["dst_width"][0]), int(karg["dst_height"][0])
    tr = tracer_mod.Tracer(w, h, int(karg["samples_per_pixel"][0]), int(karg["max_depth"][0]), cuda_device=cuda_device)
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
