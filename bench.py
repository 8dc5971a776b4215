#!/usr/bin/env python
"""bench.py — Mrays/s of the ray-tracing hot path (BVH traversal + Moller-Trumbore + closest-hit shading) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One "step" = one frame of the workload through vx_start (+ the NCCL framebuffer gather when N > 1).  Rank 0 prints ONE
JSON line.  `value` = rays of all ranks / max-over-ranks device time, scene and framebuffer resident in HBM; `e2e` = the same
frame through the reference-facing call sequence of the raycast host (vx_copy_to_dev of every scene buffer from pinned host
memory, vx_upload_bytes(kernel_arg_t), vx_start, vx_ready_wait, vx_copy_from_dev of the framebuffer) timed on the host clock.
`--impl reference` times the reference's own CPU implementation of the same frame (oracle/_ref = the reference sources
compiled in place; the C port when that library did not travel) on all host threads.
Workloads are BASELINE.json's configs; rays = every TLASIntersect-equivalent query (primary + shadow + bounce).
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
METRIC = "Mrays/s (primary + shadow + bounce rays, whole job)"
BAND_ROWS = 8


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


def gather_roofline(dev, achieved_gbs, beyond_l2=False):
    """second roofline denominator: random-record gather bandwidth into the SMs, measured live with the library's
    probe kernel (vx_b200_probe_gather): 32-byte records = one reference box, 128-byte = one 4-wide record of the fast order.
    The scratch buffer is L2-resident (32 MB) — or 1 GB, when the workload's scene is far beyond L2."""
    size = (1 << 30) if beyond_l2 else (32 << 20)
    p32, p128 = dev.probe_gather(size, 32), dev.probe_gather(size, 128)
    return {"buffer_bytes": size, "resident_in": "HBM (buffer >> L2)" if beyond_l2 else "L2", "peak_32B_records": p32, "peak_128B_records": p128,
            "unit": "GB/s", "achieved": achieved_gbs, "frac": achieved_gbs / p32 if p32 > 0 else None,
            "note": "achieved = the same algorithmic bytes/s as roofline.achieved, against the 32-byte-record gather rate"}


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(p.get("sm_max_mhz", 1965.0))
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled while the timed region runs (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
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
    steps instead of simulated cycles: a step = one node fetch + its box tests, or one leaf."""
    if not st.get("warp_steps"):
        return None
    return {"order": order, "warps": st["warps"], "warp_steps": st["warp_steps"], "thread_steps": st["lane_steps"],
            "simt_efficiency": st["lane_steps"] / (32.0 * st["warp_steps"]),
            "thread_steps_per_ray": st["lane_steps"] / max(1, st["rays"])}


def algorithmic_work(st, pixels):
    """bytes and flops of the REFERENCE algorithm for the counted work (layout A)."""
    b = 32 * st["box_tests"] + 40 * st["tri_tests"] + 160 * st["blas_entered"] + (60 + 160 + 4) * st["shaded_hits"] + 4 * pixels
    f = 27 * st["box_tests"] + 53 * st["tri_tests"] + 30 * st["blas_entered"] + 80 * st["shaded_hits"]
    return b, f


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


def cpu_frame_timer(wl, budget_s):
    """returns (fn() -> (seconds, rays), description).  The sample is the whole frame unless one frame would blow the budget,
    in which case a centred band of rows (multiple of 8) is timed instead."""
    orc, kind = cpu_oracle()
    sc = orc.scene_create(make_meshes(wl))
    sc.setup(wl["w"], wl["h"], spp=1, max_depth=wl["depth"])
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
    if "rays_per_gpu" in wl:      # incoherent-ray workload: the reference's TLASIntersect over a bounded sample of the same ray set
        orc, kind = cpu_oracle()
        sc = orc.scene_create(make_meshes(wl))
        sc.setup(wl["w"], wl["h"], spp=1, max_depth=1)
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
        frame, kind, threads, sample = cpu_frame_timer(wl, budget_s=max(2.0, 150.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        frame()
    tot_t, tot_r = 0.0, 0
    for _ in range(args.steps):
        dt, n = frame()
        tot_t += dt
        tot_r += n
    v = tot_r / tot_t / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * tot_t / max(1, args.steps), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "width": wl["w"], "height": wl["h"], "triangles": 2 * wl["n"] ** 2 if wl["n"] else 32,
                       "mode": wl["mode"], "max_depth": wl["depth"]},
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# the b200 arm
# ------------------------------------------------------------------------------------------------------------------
def run_b200(args, wl, name):
    import torch
    import torch.distributed as dist

    vx, tracer_mod, host, mg = pkg("vx"), pkg("tracer"), pkg("host"), pkg("multigpu")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: the b200 backend has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H = wl["w"], wl["h"]

    # ---- host side: scene preparation with the product's own builders (reference byte layouts) ----
    t0 = time.perf_counter()
    hs = host.HostScene(make_meshes(wl))
    karg = hs.setup(W, H, spp=1, max_depth=wl["depth"])
    bufs = hs.buffers()
    build_s = time.perf_counter() - t0

    tr = tracer_mod.Tracer(W, H, 1, wl["depth"], cuda_device=local).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    tr.dev.set_partition(rank, world, BAND_ROWS)
    tr.set_mode(wl["mode"], 0, kernel=args.kernel)
    tr._upload_args()
    ext = torch.cuda.ExternalStream(tr.dev.stream(), device=torch.device("cuda", local))
    fb = mg.framebuffer_tensor(tr, local)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=f"cuda:{local}")       # > 126 MB L2

    # N > 1: by default the ranks' kernels store their bands straight into rank 0's framebuffer over NVLink peer memory
    # (--gather p2p); --gather nccl renders locally and moves the bands with one NCCL gather inside the timed region.
    p2p = world > 1 and args.gather == "p2p"
    check = {}
    if world > 1:   # both transports must deliver the same frame to rank 0
        tr.dev.start(tr.krnl, tr.args)
        with torch.cuda.stream(ext):
            mg.gather_framebuffer(fb, rank, world, BAND_ROWS)
        ext.synchronize()
        tr.dev.ready_wait()
        ref_frame = fb.clone() if rank == 0 else None
        if rank == 0:
            fb.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        mg.share_framebuffer(tr, rank, world)
        tr.dev.start(tr.krnl, tr.args)
        tr.dev.ready_wait()
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            check["p2p_frame_equals_nccl_gather_frame"] = bool(torch.equal(fb, ref_frame))
        if not p2p:
            mg.unshare_framebuffer(tr, rank, world)

    def step(timed_events=None):
        with torch.cuda.stream(ext):
            flush.zero_()                                       # L2 flush between iterations, outside the timed pair
            if timed_events:
                timed_events[0].record()
            tr.dev.start(tr.krnl, tr.args)                      # vx_start: async, on the same CUDA stream
            if world > 1 and not p2p:
                mg.gather_framebuffer(fb, rank, world, BAND_ROWS)
            if timed_events:
                timed_events[1].record()

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    sync_all()
    tr.dev.ready_wait()
    sampler = ClockSampler(local) if rank == 0 else None
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms = []
    for e in evs:
        step(e)
        tr.dev.ready_wait()
        kernel_ms.append(tr.dev.last_kernel_ms())
    sync_all()
    clocks = sampler.stop() if sampler else None
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    st = tr.dev.stats()
    rays_rank = st["rays"]
    launches_per_step = st["kernel_launches"]
    t_ms = torch.tensor([total_ms], dtype=torch.float64, device=f"cuda:{local}")
    r_cnt = torch.tensor([rays_rank], dtype=torch.int64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(r_cnt, op=dist.ReduceOp.SUM)
    total_ms, rays_step = float(t_ms.item()), int(r_cnt.item())
    value = rays_step * args.steps / (total_ms * 1e3)

    # ---- e2e: the raycast host's call sequence with HOST buffers, every step ----
    pinned = {n: torch.from_numpy(bufs[n].copy()).pin_memory() for n in tracer_mod.SCENE_BUFFERS if bufs[n].nbytes}
    out_pinned = torch.empty((H, W), dtype=torch.int32).pin_memory()
    h2d = sum(t.numel() for t in pinned.values()) + 192
    d2h = W * H * 4 if rank == 0 else 0

    # N > 1: the scene enters the box once — rank 0 uploads it from its host buffers and the other ranks receive it over NVLink
    # (NCCL broadcast into their own vx buffers + vx_b200_mark_written) instead of N uploads through the host's memory system
    bcast = world > 1 and (args.scene_upload == "broadcast" or (args.scene_upload == "auto" and world >= 4))
    scene_t = mg.scene_tensors(tr, local) if bcast else None

    def e2e_step():
        if not bcast or rank == 0:
            for n, t in pinned.items():
                tr.bufs[n].copy_to_dev((t.data_ptr(), t.numel()))
        if bcast:
            with torch.cuda.stream(ext):
                mg.broadcast_scene(tr, scene_t, rank, world)
            ext.synchronize()
        tr._upload_args()
        tr.dev.start(tr.krnl, tr.args)
        if world > 1 and not p2p:
            with torch.cuda.stream(ext):
                mg.gather_framebuffer(fb, rank, world, BAND_ROWS)
            ext.synchronize()
        tr.dev.ready_wait()
        if p2p:
            dist.barrier()          # every rank's bands have landed in rank 0's framebuffer
        if rank == 0:
            tr.out.copy_from_dev((out_pinned.data_ptr(), W * H * 4))

    # warm-up of the end-to-end sequence: the first host->device copies of a process run at ~14 GB/s and reach the link rate
    # (~50 GB/s) only after the GPU has seen some tens of ms of work (tools/quick_upload_probe2.py, profiles/r01_upload_ramp.txt),
    # so the untimed steps cover that ramp: at least W steps and at least 0.15 s, the same count on every rank
    e2e_steps = max(3, min(args.steps, 20))
    e2e_warm = max(3, args.warmup)
    t_w = time.perf_counter()
    for _ in range(e2e_warm):
        e2e_step()
    sync_all()
    per_step = max((time.perf_counter() - t_w) / e2e_warm, 1e-4)
    more = torch.tensor([min(200, max(0, int(0.15 / per_step) - e2e_warm))], dtype=torch.int64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(more, op=dist.ReduceOp.MAX)
    for _ in range(int(more.item())):
        e2e_step()
    e2e_warm += int(more.item())
    sync_all()
    e2e_batches = []            # three timed batches of e2e_steps frames; the line carries the median batch (host timing jitters)
    for _ in range(3):
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        sync_all()
        e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=f"cuda:{local}")
        if world > 1:
            dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
        e2e_batches.append(rays_step * e2e_steps / float(e2e_s.item()) / 1e6)
    e2e_value = sorted(e2e_batches)[1]
    e2e_launches = tr.dev.stats()["kernel_launches"]
    if world > 1 and rank == 0:   # the frame that came out of the end-to-end sequence is the frame of the transport check
        check["e2e_frame_equals_checked_frame"] = bool(np.array_equal(out_pinned.numpy().view(np.uint32), ref_frame.cpu().numpy().view(np.uint32)))
        check["scene_replication"] = ("rank 0 uploads, NCCL broadcast over NVLink to the other ranks' vx buffers" if bcast
                                      else "every rank uploads its own copy")

    # ---- roofline of the dominant kernel (k_render_a): algorithmic bytes of the reference algorithm / kernel time ----
    tr.set_mode(wl["mode"], 1)              # strict order = the reference's own visiting order => its algorithmic work
    tr.dev.enable_stats(True)
    tr._upload_args()
    tr.launch()
    work = tr.dev.stats()
    tr.set_mode(wl["mode"], 0)
    tr.launch()
    unit_perf = [rt_unit_perf(tr.dev.stats(), "fast"), rt_unit_perf(work, "strict")]
    tr.dev.enable_stats(False)
    pixels_rank = len(mg.owned_rows(rank, world, H, BAND_ROWS)) * W
    abytes, aflops = algorithmic_work(work, pixels_rank)
    k_ms = float(np.mean(kernel_ms))
    hbm, hbm_src, sm_mhz = peaks()
    achieved = abytes / (k_ms * 1e-3) / 1e9
    nsm = tr.dev.caps(0x3)
    fp32_peak = nsm * 128 * 2 * sm_mhz * 1e6 / 1e12
    gather = gather_roofline(tr.dev, achieved) if rank == 0 else None
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(name)
        except Exception:
            traffic = None

    extra = {}
    if rank == 0 and world == 1 and args.extra:
        tr.close()
        tr = None
        for other in ("primary_4k_1m", "cfg2_4k_diffuse4_1m", "cfg3_8k_primary+shadow_1m"):
            if other != name:
                extra[other] = quick_run(other, local, kernel=args.kernel)
        if "cfg3_8k_primary+shadow_1m" in extra:   # --gpus N > 1 runs BASELINE configs[3] (strong scaling); this is its one-GPU point
            extra["cfg3_8k_primary+shadow_1m"]["note"] += "; the one-GPU base of the strong-scaling series that --gpus 2/4/8 measure"
        extra["rtu_1080p_primary_100k"] = quick_run_rtu(local, with_cpu=not args.no_cpu)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        frame, kind, threads, sample = cpu_frame_timer(wl, budget_s=20.0)
        best = None
        for _ in range(2):
            dt, n = frame()
            best = (dt, n) if best is None or dt < best[0] else best
        cpu = {"value": best[1] / best[0] / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample + ", best of 2"}
        cpu["single_thread"] = frame.single()

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": name, "baseline_config_index": wl["config_index"], "width": W, "height": H,
                           "triangles": 2 * wl["n"] ** 2 if wl["n"] else 32, "mode": wl["mode"], "max_depth": wl["depth"],
                           "rays_per_step": rays_step, "partition": f"{BAND_ROWS}-row bands round-robin over {world} GPU(s), scene replicated",
                           "schedule_feedback": "on: tile groups handed out in descending order of the cost recorded on earlier frames when "
                                                "a timed trial during warm-up found that faster (DESIGN 5; VORTEX_B200_SCHEDULE=0 disables)",
                           "framebuffer_transport": ("none (single GPU)" if world == 1 else
                                                     "kernels store into rank 0's framebuffer over NVLink peer memory (CUDA IPC)" if p2p else
                                                     "NCCL gather inside the timed region"), **check,
                           "l2": "512 MB memset between timed iterations (L2 flush); timed per step with CUDA events on the launch stream",
                           "host_bvh_build_s": round(build_s, 3), "extra_workloads": extra},
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps, "warmup": e2e_warm, "batches": e2e_batches,
                        "timer": "host wall clock around the synchronous vx_* calls, max over ranks", "launches_per_step": e2e_launches},
                "gpu_launches": int(launches_per_step) * args.steps, "rt_unit_perf": unit_perf,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": traffic,
                             "peak_source": hbm_src, "kernel": "k_render_a" if args.kernel == 1 else "k_render_m", "kernel_ms": k_ms,
                             "algorithmic_bytes_per_launch": abytes, "algorithmic_bytes_per_ray": abytes / max(1, work["rays"]),
                             "fp32": {"achieved_tflops": aflops / (k_ms * 1e-3) / 1e12, "peak_tflops": fp32_peak,
                                      "frac": aflops / (k_ms * 1e-3) / 1e12 / fp32_peak},
                             "work_per_ray": {k: work[k] / max(1, work["rays"]) for k in ("box_tests", "tri_tests", "blas_entered", "shaded_hits")},
                             "l2_gather": gather,
                             "note": "achieved = the REFERENCE algorithm's bytes for the counted work (strict-order counters) / kernel time, as "
                                     "the contract defines it; those bytes are served from L1/L2 (`traffic` is what reached DRAM) and the "
                                     "4-wide records read fewer of them, so frac can exceed 1 — the binding resource is instruction issue "
                                     "(ncu: 80 % issue-active, profiles/r01_ncu_machine_cfg1_sched.txt)"},
                "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    if tr is not None:
        tr.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_incoherent(args, wl, name):
    """BASELINE configs[4]: closest-hit records of incoherent rays through vx_b200_trace_rays.  Rays are sharded by
    index (rank r traces rays [r*R, (r+1)*R) of one global WangHash-seeded set), the scene is replicated, no data-path collective:
    weak scaling.  value = rays of all ranks / max-over-ranks device time with rays and scene resident in HBM; e2e adds the
    host->device copy of the rays and the device->host copy of the hit records every step."""
    import torch
    import torch.distributed as dist

    vx, tracer_mod, host, scenes = pkg("vx"), pkg("tracer"), pkg("host"), pkg("scenes")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: the b200 backend has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    R = int(wl["rays_per_gpu"])
    t0 = time.perf_counter()
    hs = host.HostScene(make_meshes(wl))
    karg = hs.setup(wl["w"], wl["h"], spp=1, max_depth=1)
    bufs = hs.buffers()
    build_s = time.perf_counter() - t0
    tr = tracer_mod.Tracer(wl["w"], wl["h"], 1, 1, cuda_device=local).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    tr.set_mode(0, 0)
    tr._upload_args()
    sc_ = np.float32(wl.get("scale", 1.0))
    rays = scenes.random_rays(R, tuple(sc_ * np.float32(v) for v in (-1.0, -0.05, -1.0)), tuple(sc_ * np.float32(v) for v in (1.0, 0.45, 1.0)),
                              seed=1, start=rank * R)   # the scene's AABB, a little headroom above
    rays_pinned = torch.from_numpy(rays).pin_memory()
    hits_pinned = torch.empty(R * 24, dtype=torch.uint8).pin_memory()
    rb = tr.dev.mem_alloc(R * 24)
    hb = tr.dev.mem_alloc(R * 24)
    rb.copy_to_dev((rays_pinned.data_ptr(), R * 24))
    ext = torch.cuda.ExternalStream(tr.dev.stream(), device=torch.device("cuda", local))
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=f"cuda:{local}")

    def step(ev=None):
        with torch.cuda.stream(ext):
            flush.zero_()
            if ev:
                ev[0].record()
            tr.dev.trace_rays(tr.args, rb, R, hb)
            if ev:
                ev[1].record()

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    sync_all()
    tr.dev.ready_wait()
    sampler = ClockSampler(local) if rank == 0 else None
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms = []
    for e in evs:
        step(e)
        tr.dev.ready_wait()
        kernel_ms.append(tr.dev.last_kernel_ms())
    sync_all()
    clocks = sampler.stop() if sampler else None
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    launches_per_step = tr.dev.stats()["kernel_launches"]
    t_ms = torch.tensor([total_ms], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    total_ms = float(t_ms.item())
    rays_step = R * world
    value = rays_step * args.steps / (total_ms * 1e3)

    def e2e_step_calls():      # the four reference-shaped calls, one after the other
        rb.copy_to_dev((rays_pinned.data_ptr(), R * 24))
        tr.dev.trace_rays(tr.args, rb, R, hb)
        tr.dev.ready_wait()
        hb.copy_from_dev((hits_pinned.data_ptr(), R * 24))

    def e2e_step():            # the host-buffer entry point: same records, upload / traversal / download of chunks overlapped
        tr.dev.trace_rays_host(tr.args, rays_pinned.data_ptr(), R, hits_pinned.data_ptr())

    def e2e_measure(fn, n):
        fn()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        sync_all()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=f"cuda:{local}")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        return rays_step * n / float(dt.item()) / 1e6

    e2e_steps = max(3, min(args.steps, 5))
    e2e_calls_value = e2e_measure(e2e_step_calls, e2e_steps)
    calls_bytes = hits_pinned.numpy().tobytes()
    hits_pinned.zero_()
    e2e_value = e2e_measure(e2e_step, e2e_steps)
    e2e_same = hits_pinned.numpy().tobytes() == calls_bytes
    host_launches = tr.dev.stats()["kernel_launches"]
    hits = np.frombuffer(hits_pinned.numpy().tobytes(), dtype=tracer_mod.HIT_DT)
    hit_frac = float((hits["dist"] < 1e30).mean())

    tr.set_mode(0, 1)                      # strict order = the reference's visiting order => its algorithmic work
    tr.dev.enable_stats(True)
    tr.dev.trace_rays(tr.args, rb, R, hb)
    tr.dev.ready_wait()
    work = tr.dev.stats()
    tr.set_mode(0, 0)
    tr.dev.trace_rays(tr.args, rb, R, hb)
    tr.dev.ready_wait()
    unit_perf = [rt_unit_perf(tr.dev.stats(), "fast"), rt_unit_perf(work, "strict")]
    tr.dev.enable_stats(False)
    abytes = 32 * work["box_tests"] + 40 * work["tri_tests"] + 160 * work["blas_entered"] + 48 * R    # + ray in (24 B) and hit record out (24 B)
    aflops = 27 * work["box_tests"] + 53 * work["tri_tests"] + 30 * work["blas_entered"]
    k_ms = float(np.mean(kernel_ms))
    hbm, hbm_src, sm_mhz = peaks()
    achieved = abytes / (k_ms * 1e-3) / 1e9
    fp32_peak = tr.dev.caps(0x3) * 128 * 2 * sm_mhz * 1e6 / 1e12
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(name)
    except Exception:
        pass

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        po = importlib.import_module("oracle.pyoracle")
        orc, kind = cpu_oracle()
        raw = po.RawScene(orc, bufs, karg)       # the oracle's hot loop on the very bytes that were uploaded
        sample = rays[: 1 << 18]
        threads = orc.hw_threads()
        best = None
        for _ in range(2):
            t0 = time.perf_counter()
            ref_hits = raw.trace_rays(sample, threads=threads)
            dt = time.perf_counter() - t0
            best = dt if best is None or dt < best else best
        same = (ref_hits["triIdx"] == hits["triIdx"][: len(sample)]) & (ref_hits["dist"] == hits["dist"][: len(sample)])
        cpu = {"value": len(sample) / best / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
               "sample": f"first {len(sample)} rays of rank 0's shard, all host threads, best of 2",
               "gpu_hit_records_identical_on_sample": float(same.mean())}
    if rank == 0:
        ntri = 2 * wl["n"] ** 2
        line = {"metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": name, "baseline_config_index": wl["config_index"], "triangles": ntri, "scene_scale": wl.get("scale", 1.0), "rays_per_gpu": R,
                           "rays_per_step": rays_step, "hit_fraction": hit_frac,
                           "partition": f"rays sharded by index over {world} GPU(s), scene replicated, no collective on the data path",
                           "l2": "512 MB memset between timed iterations (L2 flush); working set (scene 1.6 GB at 10 M triangles) >> L2",
                           "host_bvh_build_s": round(build_s, 3)},
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": R * 24, "d2h_bytes_per_step": R * 24, "steps": e2e_steps,
                        "timer": "host wall clock around vx_b200_trace_rays_host (pinned host rays in, pinned host hit records out; "
                                 "chunks pipelined on three streams), max over ranks",
                        "launches_per_step": int(host_launches),
                        "four_call_sequence": {"value": e2e_calls_value, "unit": "Mrays/s", "identical_records": bool(e2e_same),
                                               "calls": "vx_copy_to_dev(rays) + vx_b200_trace_rays + vx_ready_wait + vx_copy_from_dev(hits)"}},
                "gpu_launches": int(launches_per_step) * args.steps, "rt_unit_perf": unit_perf,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": traffic,
                             "peak_source": hbm_src, "kernel": "k_render_m<trace> (ray state machine, rays as work items)", "kernel_ms": k_ms, "algorithmic_bytes_per_launch": abytes,
                             "algorithmic_bytes_per_ray": abytes / R,
                             "fp32": {"achieved_tflops": aflops / (k_ms * 1e-3) / 1e12, "peak_tflops": fp32_peak,
                                      "frac": aflops / (k_ms * 1e-3) / 1e12 / fp32_peak},
                             "work_per_ray": {k: work[k] / R for k in ("box_tests", "tri_tests", "blas_entered")},
                             "l2_gather": gather_roofline(tr.dev, achieved, beyond_l2=ntri >= 4_000_000)},
                "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    tr.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def quick_run(name, local, steps=5, kernel=0):
    """kernel-only Mrays/s of another workload on one GPU (reported under config.extra_workloads; not the headline)."""
    tracer_mod, host = pkg("tracer"), pkg("host")
    wl = WORKLOADS[name]
    hs = host.HostScene(make_meshes(wl))
    karg = hs.setup(wl["w"], wl["h"], spp=1, max_depth=wl["depth"])
    tr = tracer_mod.Tracer(wl["w"], wl["h"], 1, wl["depth"], cuda_device=local).init(hs.buffers())
    tr.setup({k: karg[k][0] for k in karg.dtype.names})
    tr.set_mode(wl["mode"], 0, kernel=kernel)
    tr._upload_args()
    ms = []
    for i in range(steps + 3):
        tr.launch()
        if i >= 3:
            ms.append(tr.dev.last_kernel_ms())
    rays = tr.dev.stats()["rays"]
    tr.close()
    return {"mrays_per_s": rays / (float(np.mean(ms)) * 1e3), "kernel_ms": float(np.mean(ms)), "rays_per_step": rays,
            "note": "kernel time only, L2 not flushed (working set > L2 for the 1M-triangle scenes' framebuffer+scene)"}


def quick_run_rtu(local, n=224, w=1920, h=1080, steps=5, with_cpu=True):
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
    tr = rtx.RtxTracer(w, h, 1, 1, cuda_device=local).init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names}, camera=rtx.CAMERA_PINHOLE)
    tr._upload_args()
    ms = []
    for i in range(steps + 3):
        tr.launch()
        if i >= 3:
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
    # the same frame through the traceRay/getWork/getAttr/commit shim with the shipped shaders dispatched by SBT
    tr.dev.enable_stats(False)
    tr.set_program(0)
    pms = []
    for i in range(steps + 2):
        tr.launch()
        if i >= 2:
            pms.append(tr.dev.last_kernel_ms())
    tr.set_program(None)
    program_ms = float(np.mean(pms))
    tr.close()
    abytes = 52 * wk["box_tests"] + 52 * wk["blas_entered"] + 36 * wk["tri_tests"] + 156 * wk["shaded_hits"] + 4 * w * h
    k_ms = float(np.mean(ms))
    hbm, _, _ = peaks()
    out = {"mrays_per_s": rays / (k_ms * 1e3), "kernel": "k_rtu_render (fast order)", "kernel_ms": k_ms, "rays_per_step": rays,
           "strict_order_kernel_ms": strict_ms, "host_bvh4_build_s": round(build_s, 3),
           "algorithmic_bytes_per_ray": abytes / max(1, rays), "roofline_hbm_frac": abytes / (k_ms * 1e-3) / 1e9 / hbm,
           "work_per_ray": {"nodes": wk["box_tests"] / max(1, rays), "tri_tests": wk["tri_tests"] / max(1, rays),
                            "blas_entered": wk["blas_entered"] / max(1, rays), "shaded_hits": wk["shaded_hits"] / max(1, rays)},
           "rt_unit_perf": unit_perf,
           "shader_program_kernel": {"kernel": "k_rtu_program<ShippedShaders> (fast order)", "kernel_ms": program_ms,
                                     "mrays_per_s": rays / (program_ms * 1e3)},
           "note": "kernel time only, L2 not flushed; algorithmic work = the reference's trail/short-stack order (strict)"}
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--scene-upload", default="auto", choices=["auto", "broadcast", "replicate"],
                    help="N > 1, end-to-end leg: rank 0 uploads + NCCL broadcast, or every rank uploads; auto = broadcast from 4 GPUs up "
                         "(measured: 2 GPUs 4915 vs 5152, 4 GPUs 5757 vs 5399, 8 GPUs 6136 vs 4058 Mrays/s)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--kernel", type=int, default=0, help="layout-A render kernel: 0 = ray state machine (default), 1 = one pixel per lane")
    ap.add_argument("--gather", default="p2p", choices=["p2p", "nccl"], help="N > 1: how the bands reach rank 0's framebuffer")
    ap.add_argument("--extra", type=int, default=1, help="also report kernel-only numbers of the 1M-triangle workloads (N=1)")
    args = ap.parse_args()
    name = args.workload or ("cfg1_1080p_primary+shadow_100k" if args.gpus == 1 else "cfg3_8k_primary+shadow_1m")
    wl = WORKLOADS[name]
    if args.impl == "reference":
        run_reference(args, wl, name)
    elif "rays_per_gpu" in wl:
        run_incoherent(args, wl, name)
    else:
        run_b200(args, wl, name)


if __name__ == "__main__":
    main()
