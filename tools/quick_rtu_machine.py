"""layout B: ray state machine vs one ray per lane — frames (1080p / 100 K, 4K / 1 M) and incoherent hit records (1 M triangles,
2^22 random rays), with a sweep of the machine's scheduling quorums (scratch tool): python tools/quick_rtu_machine.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

host, rtx, scenes = bench.pkg("host"), bench.pkg("rtx_tracer"), bench.pkg("scenes")


def scene(n, w, h):
    mesh = scenes.heightfield(n)
    bmesh = {"tris": mesh["tris"], "triex": mesh["triex"], "texid": np.zeros(mesh["tris"].shape[0], np.uint32),
             "materials": [{"tex": mesh["tex"]}], "reflectivity": 0.0}
    hs = host.RtxHostScene([bmesh])
    ax, ay = np.float32(-np.pi / 4), np.float32(np.pi / 4)
    rx = np.array([[1, 0, 0, 0], [0, np.cos(ax), -np.sin(ax), 0], [0, np.sin(ax), np.cos(ax), 0], [0, 0, 0, 1]], dtype=np.float32)
    ry = np.array([[np.cos(ay), 0, np.sin(ay), 0], [0, 1, 0, 0], [-np.sin(ay), 0, np.cos(ay), 0], [0, 0, 0, 1]], dtype=np.float32)
    hs.apply_transform(rx @ ry)
    karg = hs.setup(w, h, spp=1, max_depth=1)
    tr = rtx.RtxTracer(w, h, 1, 1, cuda_device=0).init(hs.buffers())
    tr.setup({k: karg[k][0] for k in karg.dtype.names}, camera=rtx.CAMERA_PINHOLE)
    tr._upload_args()
    return tr


def frame_ms(tr, n=12):
    ms = []
    for i in range(n + 6):
        tr.dev.flush_l2()
        tr.launch()
        if i >= 6:
            ms.append(tr.dev.last_kernel_ms())
    return float(np.mean(ms)), tr.dev.stats()["rays"]


for n, w, h in ((224, 1920, 1080), (708, 3840, 2160)):
    tr = scene(n, w, h)
    os.environ["VORTEX_B200_RTU_MACHINE"] = "0"
    os.environ["VORTEX_B200_RTU_DIRECT"] = "0"
    ms, rays = frame_ms(tr)
    print(f"{w}x{h} {2 * n * n} tris: one ray per lane {rays / ms / 1e3:.0f} Mrays/s", flush=True)
    os.environ["VORTEX_B200_RTU_DIRECT"] = "1"
    ms, rays = frame_ms(tr)
    print(f"{w}x{h} {2 * n * n} tris: one ray per lane over cursor-carrying records (k_rtu_direct) {rays / ms / 1e3:.0f} Mrays/s", flush=True)
    os.environ["VORTEX_B200_RTU_MACHINE"] = "2"      # 2 = frames through the state machine too (default: hit records only)
    for refill, inner in ((12, 8), (8, 4), (16, 8), (16, 12), (24, 12), (24, 16), (28, 20), (32, 24)):
        os.environ["VORTEX_B200_RTU_QUORUM_REFILL"], os.environ["VORTEX_B200_RTU_QUORUM_INNER"] = str(refill), str(inner)
        ms, rays = frame_ms(tr)
        print(f"    machine refill {refill} inner {inner}: {rays / ms / 1e3:.0f} Mrays/s", flush=True)
    os.environ.pop("VORTEX_B200_RTU_QUORUM_REFILL"), os.environ.pop("VORTEX_B200_RTU_QUORUM_INNER")
    if n == 708:   # incoherent rays on the 1 M-triangle scene
        R = 1 << 22
        rays_h = scenes.random_rays(R, (-1.0, -0.05, -1.0), (1.0, 0.45, 1.0), seed=1)
        rb, hb = tr.dev.upload_bytes(rays_h), tr.dev.mem_alloc(R * 24)
        out = {}
        for mach in ("0", "1"):
            os.environ["VORTEX_B200_RTU_MACHINE"] = mach
            ms = []
            for i in range(8):
                tr.dev.flush_l2()
                tr.dev.trace_rays(tr.args, rb, R, hb)
                tr.dev.ready_wait()
                if i >= 3:
                    ms.append(tr.dev.last_kernel_ms())
            a = np.zeros(R * 24, np.uint8)
            hb.copy_from_dev(a)
            out[mach] = a
            print(f"  incoherent 2^22 rays, 1 M tris: {'machine' if mach == '1' else 'one ray per lane'} {R / np.mean(ms) / 1e3:.0f} Mrays/s", flush=True)
        print("  identical records:", bool(np.array_equal(out["0"], out["1"])))
    tr.close()
