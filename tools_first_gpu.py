"""first GPU bring-up: parity vs the reference harness + rough Mrays/s (scratch script)."""
import sys, time, json
import numpy as np
sys.path.insert(0, "tests")
from rt_common import scenes, tracer_mod, vx, pyoracle, channel_diff, hit_parity

po = pyoracle()
ref = po.RefRaycast()

def run(name, meshes, w, h, spp=1, depth=1, mode=0, time_it=False, check=True):
    sc = ref.scene_create(meshes)
    karg = sc.setup(w, h, spp=spp, max_depth=depth)
    bufs = sc.buffers()
    tr = tracer_mod.Tracer(w, h, spp, depth).init(bufs)
    fields = {k: karg[k][0] for k in karg.dtype.names}
    tr.setup(fields)
    out = {"name": name, "tris": int(sum(m["tris"].shape[0] for m in meshes)), "w": w, "h": h, "mode": mode}
    for strict in (1, 0):
        tr.set_mode(mode, strict)
        img = tr.run()
        st = tr.dev.stats()
        key = "strict" if strict else "fast"
        out[key] = {"kernel_ms": st["kernel_ms"], "rays": st["rays"], "mrays": st["rays"] / (st["kernel_ms"] * 1e3) if st["kernel_ms"] else 0}
        if check:
            t0 = time.time()
            rimg, nr = sc.render(mode=mode, threads=ref.hw_threads())
            out[key]["cpu_s"] = time.time() - t0
            out[key]["img_maxdiff"] = channel_diff(img, rimg)
            out[key]["img_diff_px"] = int((img != rimg).sum())
            if mode == 0:
                gh = tr.trace_primary()
                rh = sc.trace_primary(threads=ref.hw_threads())
                out[key]["hits"] = hit_parity(gh, rh)
    if time_it:
        tr.set_mode(mode, 0)
        tr.run()
        ms = []
        for _ in range(5):
            tr.launch(); ms.append(tr.dev.stats()["kernel_ms"])
        out["fast"]["ms_list"] = ms
        tr.dev.enable_stats(True); tr.launch(); out["work"] = tr.dev.stats(); tr.dev.enable_stats(False)
    tr.close()
    print(json.dumps(out), flush=True)

run("cornell64", [scenes.cornell_box()], 64, 64)
run("cornell_mirror", [scenes.cornell_box(0.5)], 128, 128, depth=4)
run("hf32", [scenes.heightfield(32)], 256, 256)
run("hf224_1080p", [scenes.heightfield(224)], 1920, 1080, time_it=True)
run("hf224_1080p_shadow", [scenes.heightfield(224)], 1920, 1080, mode=1, time_it=True)
run("multi", [scenes.heightfield(16), scenes.icosphere((0, 0.5, 0), 0.4, 2, reflectivity=0.5), scenes.cornell_box(0.3)], 320, 240, depth=3)
run("hf708_4k", [scenes.heightfield(708)], 3840, 2160, time_it=True, check=False)
run("hf708_4k_diffuse", [scenes.heightfield(708)], 3840, 2160, depth=4, mode=2, time_it=True, check=False)
