"""Generates the RT-unit (layout B) golden fixtures tests/golden/rtu_*.npz from the REFERENCE's own code
(oracle/_ref/libref_rtu.so: sim/simx/rt_traversal.cpp compiled verbatim + the reference raytracing host's BVH4 builder /
quantizer / TLAS + its headers' shading arithmetic).  Run in the build container only:

    make -C oracle ref port && python tests/golden/make_golden_rtu.py

Each fixture holds (a) the byte buffers the reference raytracing host would upload (tri [reordered by its builder], triEx,
triIdx, mat, tex, qbvh, blas, tlas[quantized]; the unquantized bvh buffer, unused by the RT unit, is dropped), (b) its
216-byte kernel_arg_t (addresses zeroed), (c) the reference's outputs: ray_hit_t of every primary ray and the frame.
The reference BVH4 builder is not deterministic (it reads uninitialised cluster bounds, raytracing/bvh.cpp:79-86), which
is why the buffers themselves are part of the fixture.  Rays on which the compiled reference would spin 2^32 iterations
(the `size()-1` wrap, rt_traversal.cpp:81) are few in these scenes and are simply waited for.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from rt_common import pyoracle, scenes  # noqa: E402
from rtu_common import b_mesh, rtu, translate_scale  # noqa: E402

R = rtu()
ref, port = R.RtuRef(), R.RtuPort()


def dump(name, meshes, w, h, depth, camera):
    bufs, karg = ref.build_scene(meshes, w, h, max_depth=depth)
    img = R.Image(bufs, karg)
    rays = ref.primary_rays(img, camera)
    _, work, flags = port.trace_rays(img, rays, want_work=True, want_wrap_flags=True)
    print(f"{name}: {int(flags.sum())} rays reach the reference's size()-1 wrap (about 2 s each) ...", flush=True)
    hits = ref.trace_rays(img, rays, threads=os.cpu_count() or 1)
    frame = ref.render(img, camera, threads=os.cpu_count() or 1)
    k = karg.copy()
    for f in k.dtype.names:
        if f.endswith("_addr"):
            k[f] = 0
    out = {n: bufs[n] for n in R.BUFFER_NAMES}
    out["bvh"] = np.zeros(0, np.uint8)
    out.update(karg=k.view(np.uint8), image=frame, hits=hits.view(np.uint8), camera=np.uint32(camera), wrap_flags=flags,
               bvh_nodes_used=np.uint32(bufs["bvh_nodes_used"]), tlas_nodes_used=np.uint32(bufs["tlas_nodes_used"]))
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(f"{name}: {os.path.getsize(path)} bytes, hits {int((hits['dist'] < 1e30).sum())}/{len(hits)}, "
          f"frame sha1 {hashlib.sha1(frame.tobytes()).hexdigest()[:12]}, work {work}")


# 1: the reference's shipped teapot (triangles read through the reference raycast asset loader), one textured material
a = pyoracle().RefRaycast().scene_load_assets("teapot.obj", 1).buffers()
teapot = {"tris": a["tri"].view(np.float32).reshape(-1, 9), "triex": a["triEx"].view(np.float32).reshape(-1, 15),
          "tex": scenes.noise_texture(64), "reflectivity": 0.0}
dump("rtu_teapot_64", [b_mesh(teapot)], 64, 64, 1, 1)
# 2: three instances (two of them mirrors, nested traceRay recursion of closest.cpp), textured + flat materials
meshes = [b_mesh(scenes.heightfield(16), two_materials=True),
          b_mesh(scenes.icosphere((0, 0.5, 0), 0.4, 2), diffuse=(0.2, 0.6, 0.9), reflectivity=0.5),
          b_mesh(scenes.icosphere((0, 0, 0), 0.5, 2), reflectivity=0.3, transform=translate_scale((1.6, 0.6, 0.2), 1.0))]
dump("rtu_multi3_d3_96", meshes, 96, 72, 3, 1)
# 3: the camera the shipped RTU kernel hard-codes (kernel.cpp:28-39)
meshes = [b_mesh(scenes.icosphere((0, 0, 0), 1.0, 3), transform=translate_scale((40.0, 100.0, 0.0), 12.0)),
          b_mesh(scenes.heightfield(24), transform=translate_scale((60.0, 80.0, 0.0), 40.0))]
dump("rtu_fixedcam_80x48", meshes, 80, 48, 1, 0)
