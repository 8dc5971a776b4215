#!/usr/bin/env python
"""bench.py — Mrays/s of the ray-tracing hot path (BVH traversal + Moller-Trumbore + closest-hit shading) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

ONE workload for every --gpus N (1 included): BASELINE configs[3], the 7680x4320 primary + shadow frame of the 1 002 528-triangle
scene, so the N = 1/2/4/8 values form a strong-scaling curve.  One "step" = one frame through vx_start.  Multi-GPU is a
property of the BACKEND: one process opens ONE vx_device_h that spans N GPUs (vx_b200_select_gpus / env
VORTEX_B200_GPUS; slices of every upload over N PCIe links + NVLink replication, bands of the image per GPU, band-wise
download), exactly what a Vortex host application gets.  Under torchrun rank 0 drives that device; the other ranks join the
NCCL rendezvous and then wait (no GPU work of their own).  Rank 0 prints ONE JSON line:

  value     rays of the step / device time of the step — CUDA events on GPU 0's stream, which records the begin event every
            GPU waits for and joins every GPU's end event (vx_b200_stats_t.group_ms) — scene and framebuffer resident in HBM,
            L2 flushed before every step;
  e2e       the same frame through the raycast host's call sequence with HOST buffers every step: vx_copy_to_dev of the seven
            scene buffers from pinned memory, vx_upload_bytes(kernel_arg_t), vx_start, vx_ready_wait, vx_copy_from_dev of the
            framebuffer; host wall clock;
  parity    the timed frame against the CPU checker (oracle/_ref = the reference's own sources, else the C port) on the very
            bytes that were uploaded: 64 rows spread over the image pixel by pixel, and 2^18 primary rays as hit records;
  roofline  what binds the kernel according to ncu (profiles/r02_ncu_metrics.json, captures of these workloads), with the live
            kernel time: SIMT-adjusted issue utilisation, physical DRAM bytes/s against the HBM peak, algorithmic bytes of the
            reference's order and of the kernel's own order;
  config.secondary_workloads   BASELINE configs[2], the 4K primary frame, configs[1] (and configs[4] on 8 GPUs) measured the
            same way (L2 flushed, >= 20 steps, own e2e / roofline / parity).

`--impl reference` times the reference's own CPU implementation of the same frame (oracle/_ref compiled from the reference
sources; the C port when that library did not travel) on all host threads.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":   # keeps NCCL's banner off stdout: rank 0 prints ONE JSON line
    os.environ["NCCL_DEBUG"] = "WARN"

WORKLOADS = {
    # name: heightfield quads per side, width, height, mode (bit0 shadow, bit1 diffuse bounces), max_depth
    "cfg1_1080p_primary+shadow_100k": dict(n=224, w=1920, h=1080, mode=1, depth=1, config_index=1),
    "cfg2_4k_diffuse4_1m": dict(n=708, w=3840, h=2160, mode=2, depth=4, config_index=2),
    "cfg3_8k_primary+shadow_1m": dict(n=708, w=7680, h=4320, mode=1, depth=1, config_index=3),
    "primary_4k_1m": dict(n=708, w=3840, h=2160, mode=0, depth=1, config_index=None),
    "cfg0_cornell64_primary": dict(n=0, w=64, h=64, mode=0, depth=1, config_index=0),
    # BASELINE configs[4]: incoherent random-direction rays, 10 M triangles, rays sharded by index (2^23 per GPU = 2^26 on 8)
    # geometry scaled x4: at 10 M triangles over [-1,1]^2 the Moller-Trumbore determinant |a| ~ 2*area falls below the reference's
    # EPSILON = 1e-6 "parallel ray" rejection (geometry.h:1424) and 96 % of the triangles become invisible to the reference itself
    "cfg4_incoherent_10m": dict(n=2237, w=8, h=8, mode=0, depth=1, config_index=4, rays_per_gpu=1 << 23, scale=4.0),
    "incoherent_1m": dict(n=708, w=8, h=8, mode=0, depth=1, config_index=None, rays_per_gpu=1 << 23),
}
DEFAULT_WORKLOAD = "cfg3_8k_primary+shadow_1m"
SECONDARY = ("cfg2_4k_diffuse4_1m", "primary_4k_1m", "cfg1_1080p_primary+shadow_100k")
METRIC = "Mrays/s (primary + shadow + bounce rays, whole job)"
BAND_ROWS = 8
NCU_METRICS = os.path.join(ROOT, "profiles", "r02_ncu_metrics.json")


def pkg(name):
    return importlib.import_module("vortex-raytracing_b200." + name)


def make_meshes(wl):
    scenes = pkg("scenes")
    if wl["n"] == 0:
        return [scenes.cornell_box()]
    m = scenes.heightfield(wl["n"])
    if wl.get("scale", 1.0) != 1.0:
        m["tris"] = (m["tris"] * np.float32(wl["scale"])).astype(np.float32)
    return [m]


_HOST_SCENES = {}


def host_scene(wl):
    """the product's own host builders (reference byte layouts).  The procedural mesh is generated once per geometry; the scene
    (BVH2 + TLAS) is rebuilt for every workload because Tracer::setup rotates the instances in place (raycast/tracer.cpp:157-158),
    so a scene that was set up twice would show a different view."""
    key = (wl["n"], wl.get("scale", 1.0))
    if key not in _HOST_SCENES:
        _HOST_SCENES.clear()            # one 10 M-triangle mesh is 1 GB of host arrays: keep only the latest
        _HOST_SCENES[key] = make_meshes(wl)
    t0 = time.perf_counter()
    hs = pkg("host").HostScene(_HOST_SCENES[key])
    return hs, time.perf_counter() - t0


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(p.get("sm_max_mhz", 1965.0))
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


def ncu_metrics(name):
    try:
        return json.load(open(NCU_METRICS)).get(name)
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled while the timed region runs (B200_PROFILING.md recipe), all GPUs in use."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, indices):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                                          ",".join(str(i) for i in indices)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def rt_unit_perf(st, order):
    """RT-unit style counters of one run (the shape of RTUnit::PerfStats, sim/simx/rt_unit.h:15-45) in traversal
    steps / lane-iterations instead of simulated cycles."""
    if not st.get("warp_steps"):
        return None
    out = {"order": order, "warps": st["warps"], "warp_steps": st["warp_steps"], "thread_steps": st["lane_steps"],
           "simt_efficiency": st["lane_steps"] / (32.0 * st["warp_steps"]),
           "thread_steps_per_ray": st["lane_steps"] / max(1, st["rays"])}
    if any(any(r) for r in st.get("latency_dist", [])):
        out["latency_dist"] = {"rows": ["warp fetching work (GEN)", "warp shading (POST)", "warp traversing (TRAV)"],
                               "cols": ["awaiting_processing", "awaiting_scheduling", "awaiting_mf", "executing_op", "trace_complete"],
                               "lane_iterations": st["latency_dist"]}
    if any(st.get("steps_hist", [])):
        out["steps_per_ray_hist_log2"] = st["steps_hist"]
    return out


def algorithmic_work(st, pixels):
    """bytes and flops of the REFERENCE algorithm for the counted work (layout A)."""
    b = 32 * st["box_tests"] + 40 * st["tri_tests"] + 160 * st["blas_entered"] + (60 + 160 + 4) * st["shaded_hits"] + 4 * pixels
    f = 27 * st["box_tests"] + 53 * st["tri_tests"] + 30 * st["blas_entered"] + 80 * st["shaded_hits"]
    return b, f


def own_work(st, pixels, ray_io_bytes=0):
    """bytes the KERNEL's own (fast) order asks the memory system for: one 128-byte 4-wide record per four box tests, a 48-byte
    edge-form triangle per test, 48 bytes of instance transform per entry, the shading records, the pixel (or ray + hit record)"""
    return (128 * (st["box_tests"] // 4) + 48 * st["tri_tests"] + 48 * st["blas_entered"] + (60 + 160 + 4) * st["shaded_hits"]
            + 4 * pixels + ray_io_bytes)


def roofline_block(name, kernel, k_ms, n_gpus, work_ref, work_own, abytes, aflops, own_bytes, nsm_total):
    """What binds the kernel (ncu capture of THIS workload, profiles/r02_ncu_metrics.json) scaled by the live kernel time."""
    hbm, hbm_src, sm_mhz = peaks()
    m = ncu_metrics(name)
    sec = k_ms * 1e-3
    fp32_peak = nsm_total * 128 * 2 * sm_mhz * 1e6 / 1e12
    issue_peak = nsm_total * 4 * 32 * sm_mhz * 1e6 / 1e9            # G thread-instructions/s: 4 schedulers x 32 lanes per SM and clock
    out = {"kernel": kernel, "kernel_ms": k_ms, "peak_source": hbm_src,
           "algorithmic": {"reference_order_bytes_per_launch": abytes, "reference_order_gbs": abytes / sec / 1e9,
                           "kernel_order_bytes_per_launch": own_bytes, "kernel_order_gbs": own_bytes / sec / 1e9,
                           "bytes_per_ray_reference_order": abytes / max(1, work_ref["rays"]),
                           "bytes_per_ray_kernel_order": own_bytes / max(1, work_own["rays"]),
                           "note": "SURVEY 8(d) formula on the strict-order counters (= the reference's own walk) and on the fast "
                                   "order's own fetches (4-wide records); served mostly from L1/L2, so neither is an HBM fraction"},
           "fp32": {"achieved_tflops": aflops / sec / 1e12, "peak_tflops": fp32_peak, "frac": aflops / sec / 1e12 / fp32_peak},
           "work_per_ray": {"reference_order": {k: work_ref[k] / max(1, work_ref["rays"]) for k in ("box_tests", "tri_tests", "blas_entered", "shaded_hits")},
                            "kernel_order": {k: work_own[k] / max(1, work_own["rays"]) for k in ("box_tests", "tri_tests", "blas_entered", "shaded_hits")}}}
    if not m:
        out.update({"bound": "issue", "achieved": None, "peak": issue_peak, "unit": "G thread-instructions/s", "frac": None, "traffic": None,
                    "note": "no ncu capture of this workload at this commit (profiles/r02_ncu_metrics.json)"})
        return out
    g = max(1, n_gpus)
    thread_inst = m["thread_inst_executed"] / (m.get("gpus", 1)) * g if m.get("scale_with_gpus") else m["thread_inst_executed"]
    achieved = thread_inst / sec / 1e9
    dram = m["dram_bytes"]
    out.update({"bound": m.get("bound", "issue"), "achieved": achieved, "peak": issue_peak, "unit": "G thread-instructions/s",
                "frac": achieved / issue_peak, "traffic": dram,
                "hbm": {"achieved_gbs": dram / sec / 1e9, "peak_gbs": hbm * g, "frac": dram / sec / 1e9 / (hbm * g)},
                "ncu": {k: m[k] for k in m if k not in ("thread_inst_executed", "dram_bytes")}, "ncu_file": m.get("file"),
                "note": "achieved = thread-instructions per launch (ncu smsp__thread_inst_executed.sum of this workload) / live kernel "
                        "time; peak = SMs x 4 schedulers x 32 lanes x SM clock; frac = issue-slot utilisation x SIMT efficiency. "
                        "`traffic` = DRAM bytes per launch from the same capture; `hbm.frac` = those bytes / live time / measured HBM "
                        "peak.  The work per launch is fixed by the workload (total over the GPUs), the time is this run's."})
    return out


# ------------------------------------------------------------------------------------------------------------------
# CPU legs — the only places this file touches oracle/ (the checker / reported baseline, never the thing shipped)
# ------------------------------------------------------------------------------------------------------------------
def cpu_oracle():
    po = importlib.import_module("oracle.pyoracle")
    if po.have_ref():
        return po.RefRaycast(), "reference"
    if not po.have_port():
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, stdout=subprocess.DEVNULL)
    return po.PortRaycast(), "port"


def cpu_frame_timer(sc, orc, kind, wl, budget_s):
    """returns (fn() -> (seconds, rays), kind, threads, description).  The sample is the whole frame unless one frame would blow
    the budget, in which case a centred band of rows (multiple of 8) is timed instead."""
    threads = orc.hw_threads()
    H = wl["h"]
    probe_rows = (H // 2 // 8 * 8, min(H, H // 2 // 8 * 8 + 8))
    t0 = time.perf_counter()
    sc.render(mode=wl["mode"], rows=probe_rows, threads=threads)
    per_row = (time.perf_counter() - t0) / (probe_rows[1] - probe_rows[0])
    nrows = H if per_row * H <= budget_s else max(8, int(budget_s / per_row) // 8 * 8)
    y0 = (H - nrows) // 2 // 8 * 8
    rows = (y0, y0 + nrows)

    def frame():
        t = time.perf_counter()
        _, n = sc.render(mode=wl["mode"], rows=rows, threads=threads)
        dt = time.perf_counter() - t
        if wl["mode"] == 0:
            n = (rows[1] - rows[0]) * wl["w"]      # reflectivity 0 in every benchmark scene: one ray per pixel
        return dt, n

    # the reference as written is single-threaded (SURVEY.md 8d "CPU-1"): the same loop on one thread, over <= ~5 s of rows
    n1 = min(nrows, max(8, int(5.0 / (per_row * threads)) // 8 * 8))
    bands = [rows[0] + (i * (nrows // 8) // (n1 // 8)) * 8 for i in range(n1 // 8)]   # 8-row bands spread evenly over the sample's rows

    def single():
        dt, n = 0.0, 0
        for y in bands:
            t = time.perf_counter()
            _, k = sc.render(mode=wl["mode"], rows=(y, y + 8), threads=1)
            dt += time.perf_counter() - t
            n += 8 * wl["w"] if wl["mode"] == 0 else k
        return {"value": n / dt / 1e6, "unit": "Mrays/s", "cores": 1,
                "sample": f"{len(bands)} bands of 8 rows spread evenly over rows {rows[0]}..{rows[1]}, one thread (the reference as written)"}

    frame.single = single
    what = "whole frame" if nrows == H else f"rows {rows[0]}..{rows[1]} of {H}"
    return frame, kind, threads, f"{what} of {wl['w']}x{wl['h']}, all host threads"


def run_reference(args, wl, name):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    orc, kind = cpu_oracle()
    sc = orc.scene_create(make_meshes(wl))          # the reference's own Scene / BVH builders
    sc.setup(wl["w"], wl["h"], spp=1, max_depth=wl["depth"])
    if "rays_per_gpu" in wl:      # incoherent-ray workload: the reference's TLASIntersect over a bounded sample of the same ray set
        threads = orc.hw_threads()
        n_s = 1 << 18
        sc_ = np.float32(wl.get("scale", 1.0))
        rays = pkg("scenes").random_rays(n_s, tuple(sc_ * np.float32(v) for v in (-1.0, -0.05, -1.0)),
                                         tuple(sc_ * np.float32(v) for v in (1.0, 0.45, 1.0)), seed=1, start=0)

        def frame():
            t = time.perf_counter()
            sc.trace_rays(rays, threads=threads)
            return time.perf_counter() - t, n_s
        sample = f"first {n_s} rays of the global ray set, all host threads"
    else:
        frame, kind, threads, sample = cpu_frame_timer(sc, orc, kind, wl, budget_s=max(2.0, 150.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        frame()
    tot_t, tot_r = 0.0, 0
    for _ in range(args.steps):
        dt, n = frame()
        tot_t += dt
        tot_r += n
    v = tot_r / tot_t / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * tot_t / max(1, args.steps), "higher_is_better": True, "scaling": "weak" if "rays_per_gpu" in wl else "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "width": wl["w"], "height": wl["h"], "triangles": 2 * wl["n"] ** 2 if wl["n"] else 32,
                       "mode": wl["mode"], "max_depth": wl["depth"]},
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# the b200 arm
# ------------------------------------------------------------------------------------------------------------------
def sample_bands(H, n_gpus, n_bands=8):
    """8-row bands spread over the image whose indices fall on different GPUs of the round-robin deal"""
    nb = H // BAND_ROWS
    if nb <= n_bands:
        return list(range(nb))
    out = []
    for k in range(n_bands):
        b = (k * nb) // n_bands + (k % max(1, n_gpus))
        out.append(min(nb - 1, b))
    return sorted(set(out))


def parity_frames(tr, wl, bufs, karg, frame, n_gpus, max_rays=1 << 18, order="fast (the timed configuration)"):
    """the frame that was timed against the CPU checker run on the very bytes that were uploaded (raycast/tracer.cpp:249-263 is
    the loop the checker runs; north_star gates).  Pixels: 64 rows spread over the image.  Hit records: primary rays of those
    rows (at most 2^18), the same rays given to vx_b200_trace_rays and to the checker's TLASIntersect."""
    po = importlib.import_module("oracle.pyoracle")
    par = pkg("parity")
    orc, kind = cpu_oracle()
    raw = po.RawScene(orc, bufs, karg)
    threads = orc.hw_threads()
    bands = sample_bands(wl["h"], n_gpus)
    t0 = time.perf_counter()
    differ, maxd, npix = 0, 0, 0
    for b in bands:
        rows = (b * BAND_ROWS, min(wl["h"], (b + 1) * BAND_ROWS))
        ref, _ = raw.render(mode=wl["mode"], rows=rows, threads=threads)
        g, r = frame[rows[0]:rows[1]], ref[rows[0]:rows[1]]
        differ += int((g != r).sum())
        maxd = max(maxd, par.channel_diff(g, r))
        npix += g.size
    ys = [y for b in bands for y in range(b * BAND_ROWS, min(wl["h"], (b + 1) * BAND_ROWS))]
    rays = par.primary_rays(karg, ys)
    if len(rays) > max_rays:
        rays = np.ascontiguousarray(rays[:: (len(rays) + max_rays - 1) // max_rays])
    hp = par.hit_parity(tr.trace_rays(rays), raw.trace_rays(rays, threads=threads))
    return {"checker": kind, "checked_order": order, "rows": len(ys), "bands": bands, "pixels": npix,
            "pixels_differ": differ, "max_channel_diff": maxd, "rays": hp["n"], "id_mismatch": hp["mismatch"], "ties": hp["ties"],
            "non_tie": hp["non_tie"], "max_rel_t": hp["max_rel_t"], "bit_exact_t": hp["bit_exact_t"], "mismatch_frac": hp["mismatch_frac"],
            "gates": "ids equal on >= 99.99 % of rays, every residual mismatch an equal-t tie; t within 1e-5 relative; colour within 1/255",
            "green": bool(hp["non_tie"] == 0 and hp["mismatch_frac"] <= 1e-4 and hp["max_rel_t"] <= 1e-5 and maxd <= 1),
            "seconds": round(time.perf_counter() - t0, 2)}


def measure_frames(name, gpus, steps, warmup, kernel=0, cpu=False, parity=True, clocks=False):
    """one frame workload on one vx device over `gpus` GPUs: kernel-side value, e2e, roofline, parity"""
    import torch

    tracer_mod = pkg("tracer")
    wl = WORKLOADS[name]
    W, H = wl["w"], wl["h"]
    hs, build_s = host_scene(wl)
    karg = hs.setup(W, H, spp=1, max_depth=wl["depth"])
    bufs = hs.buffers()
    tr = tracer_mod.Tracer(W, H, 1, wl["depth"], cuda_device=0, gpus=gpus).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    tr.set_mode(wl["mode"], 0, kernel=kernel)
    tr._upload_args()
    dev = tr.dev
    warmup = max(4, warmup)          # the schedule trial (row-major vs recorded-cost order) takes frames 0-3

    def timed(n, warm):
        for _ in range(warm):
            dev.flush_l2()
            dev.start(tr.krnl, tr.args)
            dev.ready_wait()
        out = []
        for _ in range(n):
            dev.flush_l2()                          # L2 flush on every GPU, before the step's begin event
            dev.start(tr.krnl, tr.args)
            dev.ready_wait()
            st = dev.stats()
            out.append((st["group_ms"], st["kernel_ms"], st["kernel_ms_min"]))
        return out, dev.stats()

    # schedule feedback off = what the first frame of a new view costs (row-major tile order)
    os.environ["VORTEX_B200_SCHEDULE"] = "0"
    off, st_off = timed(max(3, min(steps, 10)), 3)
    os.environ.pop("VORTEX_B200_SCHEDULE", None)
    sampler = ClockSampler(list(range(gpus))) if clocks else None
    runs, st = timed(steps, warmup)
    clk = sampler.stop() if sampler else None
    rays_step = st["rays"]
    total_ms = sum(r[0] for r in runs)
    value = rays_step * steps / (total_ms * 1e3)
    k_ms = float(np.mean([r[1] for r in runs]))
    feedback_off = {"value": st_off["rays"] * len(off) / (sum(r[0] for r in off) * 1e3), "unit": "Mrays/s", "steps": len(off),
                    "note": "VORTEX_B200_SCHEDULE=0: tiles handed out row-major, i.e. the first frame of a new view"}

    # ---- e2e: the raycast host's call sequence with HOST buffers, every step ----
    pinned = {n: torch.from_numpy(bufs[n].copy()).pin_memory() for n in tracer_mod.SCENE_BUFFERS if bufs[n].nbytes}
    out_pinned = torch.empty((H, W), dtype=torch.int32).pin_memory()
    h2d = sum(t.numel() for t in pinned.values()) + 192
    d2h = W * H * 4

    def e2e_step():
        for n, t in pinned.items():
            tr.bufs[n].copy_to_dev((t.data_ptr(), t.numel()))
        tr._upload_args()
        dev.start(tr.krnl, tr.args)
        dev.ready_wait()
        tr.out.copy_from_dev((out_pinned.data_ptr(), d2h))

    def e2e_measure(n_steps):
        # warm-up: the first host->device copies of a process run at ~14 GB/s and reach the link rate only after the GPU has seen
        # some tens of ms of work (profiles/r01_upload_ramp.txt): at least 3 steps and 0.15 s
        t_w, n_w = time.perf_counter(), 0
        while n_w < 3 or (time.perf_counter() - t_w < 0.15 and n_w < 200):
            e2e_step()
            n_w += 1
        batches = []
        for _ in range(3):
            t0 = time.perf_counter()
            for _ in range(n_steps):
                e2e_step()
            batches.append(rays_step * n_steps / (time.perf_counter() - t0) / 1e6)
        return sorted(batches)[1], batches, n_w

    e2e_steps = max(3, min(steps, 20))
    e2e_plain, plain_batches, e2e_warm = e2e_measure(e2e_steps)
    frame_plain = out_pinned.numpy().view(np.uint32).copy()
    dev.set_readback(tr.out, out_pinned.data_ptr(), d2h)          # the image follows the kernels into host memory, band by band
    out_pinned.zero_()
    e2e_rb, rb_batches, _ = e2e_measure(e2e_steps)
    frame_rb = out_pinned.numpy().view(np.uint32).copy()
    dev.set_readback(tr.out, None, 0)
    e2e_launches = dev.stats()["kernel_launches"]
    use_rb = e2e_rb > e2e_plain
    e2e = {"value": max(e2e_rb, e2e_plain), "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
           "warmup": e2e_warm, "batches": rb_batches if use_rb else plain_batches,
           "sequence": "7 x vx_copy_to_dev (pinned host scene buffers) + vx_upload_bytes(kernel_arg_t) + vx_start + vx_ready_wait + "
                       "vx_copy_from_dev(framebuffer -> pinned host image)" + (", with vx_b200_set_readback registered once before the loop "
                       "(the framebuffer copy is issued behind the kernels on each GPU's stream)" if use_rb else ""),
           "plain_sequence": {"value": e2e_plain, "batches": plain_batches},
           "with_readback": {"value": e2e_rb, "batches": rb_batches, "same_frame": bool(np.array_equal(frame_rb, frame_plain))},
           "timer": "host wall clock around the synchronous vx_* calls", "launches_per_step": e2e_launches}

    # ---- work counters: strict order = the reference's own walk; fast order = what the timed kernel does ----
    dev.enable_stats(True)
    tr.set_mode(wl["mode"], 1)
    tr.launch()
    work_ref = dev.stats()
    tr.set_mode(wl["mode"], 0, kernel=kernel)
    tr.launch()
    work_own = dev.stats()
    dev.enable_stats(False)
    unit_perf = [rt_unit_perf(work_own, "fast"), rt_unit_perf(work_ref, "strict")]
    abytes, aflops = algorithmic_work(work_ref, W * H)
    kname = "k_render_a" if kernel == 1 else ("k_render_m" if os.environ.get("VORTEX_B200_DIRECT", "1") == "0" else "k_render_direct")
    roof = roofline_block(name, kname, k_ms, gpus, work_ref, work_own, abytes, aflops,
                          own_work(work_own, W * H), dev.caps(0x3))

    par = None
    if parity:
        tr.set_mode(wl["mode"], 0, kernel=kernel)
        frame = tr.run()
        par = parity_frames(tr, wl, bufs, karg, frame, gpus)
        par["e2e_frame_equals_timed_frame"] = bool(np.array_equal(frame_plain, frame))

    cpu_b = None
    if cpu:
        po = importlib.import_module("oracle.pyoracle")
        orc, kind = cpu_oracle()
        raw = po.RawScene(orc, bufs, karg)        # the reference's hot loop on the very bytes that were uploaded
        fr, kind, threads, sample = cpu_frame_timer(raw, orc, kind, wl, budget_s=12.0)
        best = None
        for _ in range(2):
            dt, n = fr()
            best = (dt, n) if best is None or dt < best[0] else best
        cpu_b = {"value": best[1] / best[0] / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample + ", best of 2"}
        cpu_b["single_thread"] = fr.single()
    launches = int(st["kernel_launches"])
    tr.close()
    return {"workload": name, "baseline_config_index": wl["config_index"], "value": value, "unit": "Mrays/s", "n_gpus": gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": total_ms / steps, "rays_per_step": rays_step, "width": W, "height": H,
            "triangles": 2 * wl["n"] ** 2 if wl["n"] else 32, "mode": wl["mode"], "max_depth": wl["depth"],
            "per_gpu_kernel_ms": {"slowest": k_ms, "fastest": float(np.mean([r[2] for r in runs]))},
            "schedule_feedback_off": feedback_off, "e2e": e2e, "gpu_launches": launches * steps, "rt_unit_perf": unit_perf,
            "roofline": roof, "parity": par, "cpu_baseline": cpu_b, "clocks": clk, "host_bvh_build_s": round(build_s, 3)}


def measure_incoherent(name, gpus, steps, warmup, cpu=False, clocks=False):
    """BASELINE configs[4]: closest-hit records of incoherent rays through vx_b200_trace_rays.  One global
    WangHash-seeded ray set of 2^23 rays per GPU, sharded by index over the GPUs of the device (scene replicated, no data-path
    collective: weak scaling).  e2e = vx_b200_trace_rays_host: pinned host rays in, pinned host hit records out."""
    import torch

    tracer_mod, scenes = pkg("tracer"), pkg("scenes")
    wl = WORKLOADS[name]
    R = int(wl["rays_per_gpu"]) * gpus
    # every GPU's arena mirrors the whole address space (scene + the ray and hit buffers of ALL GPUs): beyond the 4 GiB of the
    # XLEN=32 build the backend is asked for a larger arena (addresses above 4 GiB are fine for layout A's 64-bit fields)
    need_gb = (R * 48 + 170 * 2 * wl["n"] ** 2) / 2 ** 30 + 0.5
    if need_gb > 4:
        os.environ["VORTEX_B200_MEM_GB"] = str(int(need_gb) + 1)
    hs, build_s = host_scene(wl)
    karg = hs.setup(wl["w"], wl["h"], spp=1, max_depth=1)
    bufs = hs.buffers()
    tr = tracer_mod.Tracer(wl["w"], wl["h"], 1, 1, cuda_device=0, gpus=gpus).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    tr.set_mode(0, 0)
    tr._upload_args()
    dev = tr.dev
    sc_ = np.float32(wl.get("scale", 1.0))
    rays = scenes.random_rays(R, tuple(sc_ * np.float32(v) for v in (-1.0, -0.05, -1.0)), tuple(sc_ * np.float32(v) for v in (1.0, 0.45, 1.0)),
                              seed=1, start=0)   # the scene's AABB, a little headroom above
    rays_pinned = torch.from_numpy(rays).pin_memory()
    hits_pinned = torch.empty(R * 24, dtype=torch.uint8).pin_memory()
    rb = dev.mem_alloc(R * 24)
    hb = dev.mem_alloc(R * 24)
    rb.copy_to_dev((rays_pinned.data_ptr(), R * 24))
    warmup = max(3, warmup)
    for _ in range(warmup):
        dev.flush_l2()
        dev.trace_rays(tr.args, rb, R, hb)
        dev.ready_wait()
    sampler = ClockSampler(list(range(gpus))) if clocks else None
    runs = []
    for _ in range(steps):
        dev.flush_l2()
        dev.trace_rays(tr.args, rb, R, hb)
        dev.ready_wait()
        st = dev.stats()
        runs.append((st["group_ms"], st["kernel_ms"], st["kernel_ms_min"]))
    clk = sampler.stop() if sampler else None
    launches = int(dev.stats()["kernel_launches"])
    total_ms = sum(r[0] for r in runs)
    value = R * steps / (total_ms * 1e3)
    k_ms = float(np.mean([r[1] for r in runs]))

    def e2e_calls():           # the four reference-shaped calls, one after the other
        rb.copy_to_dev((rays_pinned.data_ptr(), R * 24))
        dev.trace_rays(tr.args, rb, R, hb)
        dev.ready_wait()
        hb.copy_from_dev((hits_pinned.data_ptr(), R * 24))

    def e2e_host():            # the host-buffer entry point: same records, upload / traversal / download of chunks overlapped
        dev.trace_rays_host(tr.args, rays_pinned.data_ptr(), R, hits_pinned.data_ptr())

    def e2e_measure(fn, n):
        fn()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        return R * n / (time.perf_counter() - t0) / 1e6

    e2e_steps = max(3, min(steps, 5))
    e2e_calls_value = e2e_measure(e2e_calls, e2e_steps)
    calls_bytes = hits_pinned.numpy().tobytes()
    hits_pinned.zero_()
    e2e_value = e2e_measure(e2e_host, e2e_steps)
    e2e_same = hits_pinned.numpy().tobytes() == calls_bytes
    host_launches = dev.stats()["kernel_launches"]
    hits = np.frombuffer(calls_bytes, dtype=tracer_mod.HIT_DT)
    hit_frac = float((hits["dist"] < 1e30).mean())

    dev.enable_stats(True)
    tr.set_mode(0, 1)                      # strict order = the reference's visiting order => its algorithmic work
    dev.trace_rays(tr.args, rb, R, hb)
    dev.ready_wait()
    work_ref = dev.stats()
    tr.set_mode(0, 0)
    dev.trace_rays(tr.args, rb, R, hb)
    dev.ready_wait()
    work_own = dev.stats()
    dev.enable_stats(False)
    unit_perf = [rt_unit_perf(work_own, "fast"), rt_unit_perf(work_ref, "strict")]
    abytes = 32 * work_ref["box_tests"] + 40 * work_ref["tri_tests"] + 160 * work_ref["blas_entered"] + 48 * R    # + ray in (24 B) and hit record out (24 B)
    aflops = 27 * work_ref["box_tests"] + 53 * work_ref["tri_tests"] + 30 * work_ref["blas_entered"]
    roof = roofline_block(name, "k_render_m<trace> (ray state machine, rays as work items)", k_ms, gpus, work_ref, work_own, abytes, aflops,
                          own_work(work_own, 0, 48 * R), dev.caps(0x3))

    # parity + CPU baseline: the checker's hot loop on the very bytes that were uploaded, first 2^18 rays of the set
    po = importlib.import_module("oracle.pyoracle")
    par = pkg("parity")
    orc, kind = cpu_oracle()
    raw = po.RawScene(orc, bufs, karg)
    sample = rays[: 1 << 18]
    threads = orc.hw_threads()
    best = None
    for _ in range(2 if cpu else 1):
        t0 = time.perf_counter()
        ref_hits = raw.trace_rays(sample, threads=threads)
        dt = time.perf_counter() - t0
        best = dt if best is None or dt < best else best
    hp = par.hit_parity(hits[: len(sample)], ref_hits)
    parity = {"checker": kind, "checked_order": "fast (the timed configuration)", "rays": hp["n"], "id_mismatch": hp["mismatch"], "ties": hp["ties"],
              "non_tie": hp["non_tie"], "max_rel_t": hp["max_rel_t"], "bit_exact_t": hp["bit_exact_t"],
              "green": bool(hp["non_tie"] == 0 and hp["mismatch_frac"] <= 1e-4 and hp["max_rel_t"] <= 1e-5),
              "host_entry_point_records_identical": bool(e2e_same)}
    cpu_b = {"value": len(sample) / best / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
             "sample": f"first {len(sample)} rays of the set, all host threads, best of {2 if cpu else 1}"}
    rb.free()
    hb.free()
    tr.close()
    os.environ.pop("VORTEX_B200_MEM_GB", None)
    ntri = 2 * wl["n"] ** 2
    return {"workload": name, "baseline_config_index": wl["config_index"], "value": value, "unit": "Mrays/s", "n_gpus": gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": total_ms / steps, "rays_per_step": R, "triangles": ntri, "scene_scale": wl.get("scale", 1.0),
            "hit_fraction": hit_frac, "per_gpu_kernel_ms": {"slowest": k_ms, "fastest": float(np.mean([r[2] for r in runs]))},
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": R * 24, "d2h_bytes_per_step": R * 24, "steps": e2e_steps,
                    "timer": "host wall clock around vx_b200_trace_rays_host (pinned host rays in, pinned host hit records out; chunks "
                             "pipelined on three streams per GPU)", "launches_per_step": int(host_launches),
                    "four_call_sequence": {"value": e2e_calls_value, "unit": "Mrays/s",
                                           "calls": "vx_copy_to_dev(rays) + vx_b200_trace_rays + vx_ready_wait + vx_copy_from_dev(hits)"}},
            "gpu_launches": launches * steps, "rt_unit_perf": unit_perf, "roofline": roof, "parity": parity, "cpu_baseline": cpu_b, "clocks": clk,
            "host_bvh_build_s": round(build_s, 3)}


def measure_dynamic_scene(gpus, steps, name="primary_4k_1m"):
    """SURVEY 8f-1 in use: a scene whose triangles change every frame.  Per frame, end to end with HOST buffers: upload the triangle
    and attribute buffers, build the BVH2 ON THE GPU into the reference's own node / index buffers (vx_b200_build_bvh), render,
    download the image — against the same frame with the host's binned-SAH builder (the reference's algorithm, task-parallel)
    followed by the upload of all seven buffers.  The GPU-built tree is checked with the CPU checker on the downloaded bytes."""
    import torch

    tracer_mod = pkg("tracer")
    wl = WORKLOADS[name]
    W, H = wl["w"], wl["h"]
    hs, host_build_s = host_scene(wl)
    karg = hs.setup(W, H, spp=1, max_depth=wl["depth"])
    bufs = hs.buffers()
    ntri = bufs["tri"].nbytes // 36
    tr = tracer_mod.Tracer(W, H, 1, wl["depth"], cuda_device=0, gpus=gpus).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    tr.set_mode(wl["mode"], 0)
    tr._upload_args()
    dev = tr.dev
    pinned = {n: torch.from_numpy(bufs[n].copy()).pin_memory() for n in tracer_mod.SCENE_BUFFERS if bufs[n].nbytes}
    out_pinned = torch.empty((H, W), dtype=torch.int32).pin_memory()

    def frame(gpu_build):
        for n in (("tri", "triEx") if gpu_build else pinned):
            tr.bufs[n].copy_to_dev((pinned[n].data_ptr(), pinned[n].numel()))
        if gpu_build:
            dev.build_bvh(tr.bufs["tri"], ntri, tr.bufs["bvh"], tr.bufs["triIdx"])
        tr._upload_args()
        dev.start(tr.krnl, tr.args)
        dev.ready_wait()
        tr.out.copy_from_dev((out_pinned.data_ptr(), W * H * 4))

    def timed(gpu_build, n):
        for _ in range(4):
            frame(gpu_build)
        t0 = time.perf_counter()
        for _ in range(n):
            frame(gpu_build)
        return (time.perf_counter() - t0) / n * 1e3, dev.stats()["kernel_ms"]

    host_ms, host_kernel_ms = timed(False, steps)
    frame_host_tree = out_pinned.numpy().view(np.uint32).copy()
    gpu_ms, gpu_kernel_ms = timed(True, steps)
    dev.build_bvh(tr.bufs["tri"], ntri, tr.bufs["bvh"], tr.bufs["triIdx"])
    build_ms = dev.stats()["kernel_ms"]
    frame(True)
    rays = dev.stats()["rays"]
    frame_gpu_tree = out_pinned.numpy().view(np.uint32).copy()
    # the reference's traversal on the GPU-built tree (downloaded bytes)
    gb = dict(bufs)
    for n in ("bvh", "triIdx"):
        a = np.zeros(bufs[n].nbytes, np.uint8)
        tr.bufs[n].copy_from_dev(a)
        gb[n] = a
    par = parity_frames(tr, wl, gb, karg, frame_gpu_tree, gpus, order="fast, GPU-built tree")
    par["pixels_differ_from_host_tree_frame"] = int((frame_gpu_tree != frame_host_tree).sum())
    tr.close()
    return {"workload": name + " with the triangles re-uploaded and the BVH rebuilt every frame", "triangles": ntri, "n_gpus": gpus, "steps": steps,
            "gpu_build": {"e2e_ms_per_frame": gpu_ms, "e2e_mrays_per_s": rays / gpu_ms / 1e3, "build_ms": build_ms, "render_kernel_ms": gpu_kernel_ms,
                          "sequence": "vx_copy_to_dev(tri, triEx) + vx_b200_build_bvh + vx_upload_bytes(args) + vx_start + vx_ready_wait + vx_copy_from_dev"},
            "host_build": {"e2e_ms_per_frame_excluding_build": host_ms, "host_build_ms": host_build_s * 1e3,
                           "e2e_ms_per_frame": host_ms + host_build_s * 1e3, "render_kernel_ms": host_kernel_ms,
                           "builder": "vortex-raytracing_b200/host (binned SAH, byte-identical to the reference's builder, all host threads)"},
            "tree_quality": {"render_kernel_ms_gpu_tree_over_host_tree": gpu_kernel_ms / host_kernel_ms},
            "speedup_per_frame": (host_ms + host_build_s * 1e3) / gpu_ms, "parity": par}


def quick_run_rtu(n=224, w=1920, h=1080, steps=20, with_cpu=True):
    """layout B (simx RT-unit path, quantized BVH4): kernel-only Mrays/s of k_rtu_render on the cfg1 geometry, its algorithmic
    bytes (52 B per node read + 52 B per instance entered + 36 B per triangle tested + 156 B per shaded hit + 4 B per
    pixel, counted in the reference's own visiting order) and the CPU-RTU baseline (the reference's rt_traversal.cpp when
    oracle/_ref travelled, else the C port) on a bounded band of rows."""
    host, rtx = pkg("host"), pkg("rtx_tracer")
    scenes = pkg("scenes")
    mesh = scenes.heightfield(n)
    bmesh = {"tris": mesh["tris"], "triex": mesh["triex"], "texid": np.zeros(mesh["tris"].shape[0], np.uint32),
             "materials": [{"tex": mesh["tex"]}], "reflectivity": 0.0}
    t0 = time.perf_counter()
    hs = host.RtxHostScene([bmesh])
    # the slight scene rotation the raytracing host keeps commented out (tracer.cpp:199-203) and the raycast host applies
    # (raycast/tracer.cpp:157-158): RotateX(-pi/4) * RotateY(pi/4), so that the camera looks down on the terrain as in cfg1
    ax, ay = np.float32(-np.pi / 4), np.float32(np.pi / 4)
    rx = np.array([[1, 0, 0, 0], [0, np.cos(ax), -np.sin(ax), 0], [0, np.sin(ax), np.cos(ax), 0], [0, 0, 0, 1]], dtype=np.float32)
    ry = np.array([[np.cos(ay), 0, np.sin(ay), 0], [0, 1, 0, 0], [-np.sin(ay), 0, np.cos(ay), 0], [0, 0, 0, 1]], dtype=np.float32)
    hs.apply_transform(rx @ ry)
    karg = hs.setup(w, h, spp=1, max_depth=1)
    bufs = hs.buffers()
    build_s = time.perf_counter() - t0
    tr = rtx.RtxTracer(w, h, 1, 1, cuda_device=0).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names}, camera=rtx.CAMERA_PINHOLE)
    tr._upload_args()
    ms = []
    for i in range(steps + 4):
        tr.dev.flush_l2()
        tr.launch()
        if i >= 4:
            ms.append(tr.dev.last_kernel_ms())
    rays = tr.dev.stats()["rays"]
    tr.set_order(1)
    tr.dev.enable_stats(True)
    tr.launch()
    wk = tr.dev.stats()
    strict_ms = tr.dev.last_kernel_ms()
    tr.set_order(0)
    tr.launch()
    unit_perf = [rt_unit_perf(tr.dev.stats(), "fast"), rt_unit_perf(wk, "strict")]
    tr.dev.enable_stats(False)
    # BASELINE configs[1] on this layout: primary + shadow rays (VX_DCR_B200_RT_MODE bit 0; the builder-defined extension)
    tr.set_mode(1)
    sms = []
    for i in range(steps + 4):
        tr.dev.flush_l2()
        tr.launch()
        if i >= 4:
            sms.append(tr.dev.last_kernel_ms())
    shadow_rays = tr.dev.stats()["rays"]
    tr.set_mode(0)
    # the same frame through the traceRay/getWork/getAttr/commit shim with the shipped shaders dispatched by SBT
    tr.set_program(0)
    pms = []
    for i in range(steps + 2):
        tr.dev.flush_l2()
        tr.launch()
        if i >= 2:
            pms.append(tr.dev.last_kernel_ms())
    tr.set_program(None)
    program_ms = float(np.mean(pms))
    tr.close()
    abytes = 52 * wk["box_tests"] + 52 * wk["blas_entered"] + 36 * wk["tri_tests"] + 156 * wk["shaded_hits"] + 4 * w * h
    k_ms = float(np.mean(ms))
    m = ncu_metrics("rtu_1080p_primary_100k")
    out = {"mrays_per_s": rays / (k_ms * 1e3), "kernel": "k_rtu_direct (fast order; one ray per lane over the cursor-carrying records)", "kernel_ms": k_ms, "steps": steps, "rays_per_step": rays,
           "strict_order_kernel_ms": strict_ms, "host_bvh4_build_s": round(build_s, 3),
           "algorithmic_bytes_per_ray": abytes / max(1, rays), "algorithmic_gbs_reference_order": abytes / (k_ms * 1e-3) / 1e9,
           "work_per_ray": {"nodes": wk["box_tests"] / max(1, rays), "tri_tests": wk["tri_tests"] / max(1, rays),
                            "blas_entered": wk["blas_entered"] / max(1, rays), "shaded_hits": wk["shaded_hits"] / max(1, rays)},
           "rt_unit_perf": unit_perf, "ncu": m,
           "shader_program_kernel": {"kernel": "k_rtu_program<ShippedShaders> (fast order)", "kernel_ms": program_ms,
                                     "mrays_per_s": rays / (program_ms * 1e3)},
           "primary_plus_shadow": {"kernel": "k_rtu_render<ext> (fast order)", "kernel_ms": float(np.mean(sms)), "rays_per_step": shadow_rays,
                                   "mrays_per_s": shadow_rays / (float(np.mean(sms)) * 1e3)},
           "note": "kernel time only (CUDA events), L2 flushed before every frame; algorithmic work = the reference's trail/short-stack order (strict)"}
    out["incoherent_rays"] = rtu_incoherent()
    if with_cpu:
        R = importlib.import_module("oracle.pyoracle_rtu")
        orc = R.best()
        img = R.Image(bufs, karg)
        rows = 64
        rays_cpu = orc.primary_rays(img, 1).reshape(h, w, 6)[(h - rows) // 2:(h - rows) // 2 + rows].reshape(-1, 6)
        threads = os.cpu_count() or 1
        best = None
        for _ in range(2):
            t0 = time.perf_counter()
            orc.trace_rays(img, rays_cpu, threads=threads)
            dt = time.perf_counter() - t0
            best = dt if best is None or dt < best else best
        out["cpu_rtu_baseline"] = {"value": len(rays_cpu) / best / 1e6, "unit": "Mrays/s", "cores": threads,
                                   "kind": "reference" if orc.name == "ref" else "port",
                                   "sample": f"traversal of the {rows} centre rows of {w}x{h} (sim/simx/rt_traversal.cpp order), best of 2"}
    return out


def rtu_incoherent(n=708, R=1 << 22, steps=10):
    """layout B, hit records of incoherent rays (BASELINE configs[4]'s shape on the 1 M-triangle scene): vx_b200_trace_rays through
    the ray state machine of rt_machine_b.cu against the one-ray-per-lane kernel, records checked with the CPU checker (the
    RT unit's traversal, sim/simx/rt_traversal.cpp order with the TLAS base fixed) on a sample."""
    host, rtx, scenes = pkg("host"), pkg("rtx_tracer"), pkg("scenes")
    mesh = scenes.heightfield(n)
    bmesh = {"tris": mesh["tris"], "triex": mesh["triex"], "texid": np.zeros(mesh["tris"].shape[0], np.uint32),
             "materials": [{"tex": mesh["tex"]}], "reflectivity": 0.0}
    hs = host.RtxHostScene([bmesh])
    karg = hs.setup(8, 8, spp=1, max_depth=1)
    bufs = hs.buffers()
    tr = rtx.RtxTracer(8, 8, 1, 1, cuda_device=0).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names}, camera=rtx.CAMERA_PINHOLE)
    tr._upload_args()
    rays = scenes.random_rays(R, (-1.0, -0.05, -1.0), (1.0, 0.45, 1.0), seed=1)
    rb, hb = tr.dev.upload_bytes(rays), tr.dev.mem_alloc(R * 24)
    res, recs = {}, {}
    for mach in ("0", "1"):
        os.environ["VORTEX_B200_RTU_MACHINE"] = mach
        ms = []
        for i in range(steps + 3):
            tr.dev.flush_l2()
            tr.dev.trace_rays(tr.args, rb, R, hb)
            tr.dev.ready_wait()
            if i >= 3:
                ms.append(tr.dev.last_kernel_ms())
        a = np.zeros(R * 24, np.uint8)
        hb.copy_from_dev(a)
        recs[mach] = a
        res[mach] = R / float(np.mean(ms)) / 1e3
    os.environ.pop("VORTEX_B200_RTU_MACHINE", None)
    out = {"triangles": 2 * n * n, "rays": R, "steps": steps, "mrays_per_s": res["1"], "kernel": "k_rtu_machine<trace> (ray state machine, stack ring)",
           "one_ray_per_lane_mrays_per_s": res["0"], "identical_records": bool(np.array_equal(recs["0"], recs["1"])),
           "note": "kernel time only (CUDA events), L2 flushed before every launch"}
    try:
        Rm = importlib.import_module("oracle.pyoracle_rtu")
        port = Rm.RtuPort()
        img = Rm.Image(bufs, karg)
        ns = 1 << 16
        port.set_fix_tlas_base(1)
        try:
            ref = port.trace_rays(img, rays[:ns], threads=os.cpu_count() or 1)
        finally:
            port.set_fix_tlas_base(0)
        hp = pkg("parity").hit_parity(np.frombuffer(recs["1"].tobytes(), dtype=pkg("tracer").HIT_DT)[:ns], ref)
        out["parity"] = {"checker": "port (rt_traversal.cpp restated, TLAS base fixed)", "rays": hp["n"], "id_mismatch": hp["mismatch"], "ties": hp["ties"],
                         "non_tie": hp["non_tie"], "max_rel_t": hp["max_rel_t"],
                         "green": bool(hp["non_tie"] == 0 and hp["mismatch_frac"] <= 1e-4 and hp["max_rel_t"] <= 1e-5)}
    except Exception as ex:   # the checker library did not travel
        out["parity"] = {"unavailable": str(ex)[:200]}
    rb.free()
    hb.free()
    tr.close()
    return out


def run_b200(args, name):
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    gloo = None
    if world > 1:
        import torch
        import torch.distributed as dist
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        one = torch.ones(1, device=f"cuda:{local}")
        dist.all_reduce(one)                       # the rendezvous the launcher expects: N ranks, one per GPU
        torch.cuda.synchronize()
        assert int(one.item()) == world
        gloo = dist.new_group(backend="gloo")      # the long wait below happens on the CPU, not in a spinning NCCL kernel
        if rank != 0:
            del one
            torch.cuda.empty_cache()
            dist.barrier(group=gloo)               # rank 0 drives all N GPUs through ONE vx device (the backend's multi-GPU path)
            dist.destroy_process_group()
            return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: the b200 backend has no CPU fallback")
    if torch.cuda.device_count() < args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but {torch.cuda.device_count()} visible")
    if args.fb:
        os.environ["VORTEX_B200_FB"] = args.fb
    if args.collective:
        os.environ["VORTEX_B200_COLLECTIVE"] = args.collective
    wl = WORKLOADS[name]
    incoherent = "rays_per_gpu" in wl
    head = (measure_incoherent(name, args.gpus, args.steps, args.warmup, cpu=not args.no_cpu, clocks=True) if incoherent else
            measure_frames(name, args.gpus, args.steps, args.warmup, kernel=args.kernel, cpu=not args.no_cpu, clocks=True))
    secondary = {}
    if args.extra:
        sec_steps = max(20, args.steps)
        for other in SECONDARY:
            if other != name:
                secondary[other] = measure_frames(other, args.gpus, sec_steps, args.warmup, kernel=args.kernel)
        if args.gpus >= 8 and not incoherent:
            secondary["cfg4_incoherent_10m"] = measure_incoherent("cfg4_incoherent_10m", args.gpus, sec_steps, args.warmup)
        secondary["dynamic_scene_4k_1m"] = measure_dynamic_scene(args.gpus, sec_steps)
        secondary["rtu_1080p_primary_100k"] = quick_run_rtu(steps=sec_steps, with_cpu=not args.no_cpu)
    multi = ("one vx_device_h over %d GPUs in one process (vx_b200_select_gpus): uploads sliced over %d PCIe links and replicated over "
             "NVLink (%s), %d-row bands round-robin, band-wise download (%s)" %
             (args.gpus, args.gpus, "ncclAllGather" if args.collective == "nccl" else "peer copies", BAND_ROWS,
              "kernels store into GPU 0's framebuffer over NVLink" if args.fb == "peer" else "each GPU copies its own bands to the host")
             ) if args.gpus > 1 else "single GPU"
    cfg = {"workload": name, "baseline_config_index": wl["config_index"], "triangles": head["triangles"], "rays_per_step": head["rays_per_step"],
           "multi_gpu": multi, "torchrun_world": world,
           "l2": "512 MB memset on every GPU before each timed step (vx_b200_flush_l2), outside the timed pair of events",
           "timing": "CUDA events on GPU 0's stream: begin event (all GPUs' streams wait for it) .. end event after joining every GPU's end "
                     "event (vx_b200_stats_t.group_ms)",
           "schedule_feedback": "on (DESIGN 5): tile groups handed out in descending order of the cost recorded on earlier frames when a "
                                "timed trial during warm-up found that faster; `schedule_feedback_off` carries the row-major figure",
           "host_bvh_build_s": head["host_bvh_build_s"], "secondary_workloads": secondary}
    for k in ("width", "height", "mode", "max_depth", "scene_scale", "hit_fraction"):
        if k in head:
            cfg[k] = head[k]
    line = {"metric": METRIC, "value": head["value"], "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": head["warmup"],
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak" if incoherent else "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": cfg, "clocks": head["clocks"], "e2e": head["e2e"], "gpu_launches": head["gpu_launches"],
            "parity": head["parity"], "roofline": head["roofline"], "cpu_baseline": head["cpu_baseline"], "rt_unit_perf": head["rt_unit_perf"],
            "per_gpu_kernel_ms": head["per_gpu_kernel_ms"]}
    if "schedule_feedback_off" in head:
        line["schedule_feedback_off"] = head["schedule_feedback_off"]
    print(json.dumps(line), flush=True)
    if gloo is not None:
        import torch.distributed as dist
        dist.barrier(group=gloo)
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (the parity checker still runs)")
    ap.add_argument("--kernel", type=int, default=0, help="layout-A render kernel (VX_DCR_B200_RT_KERNEL): 0 = default, 1 = round 1's one-pixel-per-lane BVH2 walk; env VORTEX_B200_DIRECT=0 = ray state machine")
    ap.add_argument("--fb", default=None, choices=["local", "peer"], help="N > 1: each GPU keeps its bands (default) / kernels store into GPU 0's framebuffer")
    ap.add_argument("--collective", default=None, choices=["peer", "nccl"], help="N > 1: scene replication by peer copies (default) / ncclAllGather")
    ap.add_argument("--extra", type=int, default=1, help="also measure the secondary workloads (config.secondary_workloads)")
    args = ap.parse_args()
    name = args.workload or DEFAULT_WORKLOAD
    if args.impl == "reference":
        run_reference(args, WORKLOADS[name], name)
    else:
        run_b200(args, name)


if __name__ == "__main__":
    main()
