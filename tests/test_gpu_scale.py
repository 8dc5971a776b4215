"""Parity at BENCHMARK scale (run on the B200 box): every BASELINE config at its real triangle count, through the same checker
block bench.py prints (bench.parity_frames: 64 rows spread over the image pixel by pixel + 2^18 primary rays as hit records,
against the CPU oracle run on the very bytes that were uploaded).  Strict order must be bit-exact, fast order within the gates.
tests/regression/raycast/render.h:143-188 (TLASIntersect) is what the checker runs per ray."""
import importlib

import numpy as np
import pytest

from rt_common import tracer_mod

pytestmark = pytest.mark.gpu
bench = importlib.import_module("bench")


def frame_case(name, w=None, h=None):
    wl = dict(bench.WORKLOADS[name])
    if w:
        wl["w"], wl["h"] = w, h
    hs, _ = bench.host_scene(wl)
    karg = hs.setup(wl["w"], wl["h"], spp=1, max_depth=wl["depth"])
    bufs = hs.buffers()
    tr = tracer_mod.Tracer(wl["w"], wl["h"], 1, wl["depth"]).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    try:
        for strict in (1, 0):
            tr.set_mode(wl["mode"], strict)
            frame = tr.run().copy()
            p = bench.parity_frames(tr, wl, bufs, karg, frame, 1, order="strict" if strict else "fast")
            assert p["green"], p
            if strict:
                assert p["pixels_differ"] == 0 and p["id_mismatch"] == 0 and p["bit_exact_t"] == p["rays"], p
            else:
                assert p["max_channel_diff"] <= 1 and p["non_tie"] == 0 and p["mismatch_frac"] <= 1e-4, p
    finally:
        tr.close()


@pytest.mark.parametrize("name,w,h", [("cfg1_1080p_primary+shadow_100k", None, None),      # BASELINE configs[1], full size
                                      ("cfg2_4k_diffuse4_1m", 960, 540),                   # configs[2]: 1 M triangles, reduced resolution
                                      ("cfg2_4k_diffuse4_1m", None, None),                 # ... and 64 rows of the full 4K frame
                                      ("cfg3_8k_primary+shadow_1m", None, None)],          # configs[3]: 64 rows of the full 8K frame
                         ids=["cfg1_full", "cfg2_1m_960x540", "cfg2_1m_4k_rows", "cfg3_1m_8k_rows"])
def test_frames_at_benchmark_scale(name, w, h):
    frame_case(name, w, h)


def test_incoherent_rays_10m_triangles():
    """BASELINE configs[4]: the 10 008 338-triangle scene (x4 scale, see bench.WORKLOADS), 2^18 random rays"""
    po = importlib.import_module("oracle.pyoracle")
    par = bench.pkg("parity")
    scenes = bench.pkg("scenes")
    wl = bench.WORKLOADS["cfg4_incoherent_10m"]
    hs, _ = bench.host_scene(wl)
    karg = hs.setup(8, 8, spp=1, max_depth=1)
    bufs = hs.buffers()
    s = np.float32(wl["scale"])
    rays = scenes.random_rays(1 << 18, tuple(s * np.float32(v) for v in (-1.0, -0.05, -1.0)), tuple(s * np.float32(v) for v in (1.0, 0.45, 1.0)), seed=1)
    orc, _ = bench.cpu_oracle()
    ref = po.RawScene(orc, bufs, karg).trace_rays(rays, threads=orc.hw_threads())
    tr = tracer_mod.Tracer(8, 8, 1, 1).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    try:
        for strict in (1, 0):
            tr.set_mode(0, strict)
            hp = par.hit_parity(tr.trace_rays(rays), ref)
            assert hp["non_tie"] == 0 and hp["mismatch_frac"] <= 1e-4 and hp["max_rel_t"] <= 1e-5, hp
            if strict:
                assert hp["mismatch"] == 0 and hp["bit_exact_t"] == hp["n"], hp
        assert (ref["dist"] < 1e30).mean() > 0.05
    finally:
        tr.close()
        bench._HOST_SCENES.clear()
