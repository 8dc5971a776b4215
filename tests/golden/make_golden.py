"""Generates the golden fixtures of tests/golden/ from the REFERENCE's own code (oracle/_ref, i.e. the reference raycast
sources compiled in place from /root/reference by oracle/Makefile).  Run in the build container only:

    make -C oracle ref && python tests/golden/make_golden.py

Each fixture is one compressed .npz holding (a) the exact byte buffers the reference host would upload (tri, triEx, triIdx,
bvh, blas, tlas, tex), (b) its kernel_arg_t (address fields zeroed), (c) the reference's outputs on them: the framebuffer of
Tracer::render and the ray_hit_t record of every primary ray.  The reference publishes no golden vectors for this path
(SURVEY.md 4); these are outputs of the reference itself run here, which is what pins the oracle port and the CUDA path.
"""
import hashlib
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from rt_common import pyoracle, scenes  # noqa: E402

po = pyoracle()
ref = po.RefRaycast()


def dump(name, sc, w, h, depth, mode=0, meshes_counts=None):
    karg = sc.setup(w, h, spp=1, max_depth=depth)
    bufs = sc.buffers()
    img, _ = sc.render(mode=mode)
    hits = sc.trace_primary()
    k = karg.copy()
    for f in k.dtype.names:
        if f.endswith("_addr"):
            k[f] = 0
    out = {n: bufs[n] for n in po.BUFFER_NAMES}
    out.update(karg=k.view(np.uint8), image=img, hits=hits.view(np.uint8), mode=np.uint32(mode),
               tri_counts=np.asarray(meshes_counts, dtype=np.uint32), bvh_nodes_used=np.uint32(bufs["bvh_nodes_used"]))
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(f"{name}: {os.path.getsize(path)} bytes, image sha1 {hashlib.sha1(img.tobytes()).hexdigest()[:12]}")


# 1-2: the reference's shipped assets through its own Mesh/Scene loaders (Tracer::init, tracer.cpp:89-100)
s = ref.scene_load_assets("teapot.obj", 1)
dump("ref_teapot_64", s, 64, 64, 1, meshes_counts=[1024])
s = ref.scene_load_assets("teapot.obj", 3)
dump("ref_teapot3_d4_64", s, 64, 64, 4, meshes_counts=[1024, 1024, 1024])
# 3: BASELINE.json configs[0] — Cornell box 64x64 primary rays
s = ref.scene_create([scenes.cornell_box()])
dump("ref_cornell_64", s, 64, 64, 1, meshes_counts=[32])
# 4: multi-instance procedural scene with mirror bounces + the shadow extension
meshes = [scenes.heightfield(16), scenes.icosphere((0, 0.5, 0), 0.4, 2, reflectivity=0.5), scenes.cornell_box(0.3)]
s = ref.scene_create(meshes)
dump("ref_multi_d3_96", s, 96, 72, 3, meshes_counts=[m["tris"].shape[0] for m in meshes])

# build self-check of the unmodified reference binary: md5 of `raycast -c -w64 -h64`
exe = os.path.join(ROOT, "oracle", "_ref", "raycast")
if os.path.exists(exe):
    subprocess.run([exe, "-c", "-w64", "-h64", "-o", "/tmp/_golden_teapot.ppm"], stdout=subprocess.DEVNULL, check=True)
    md5 = hashlib.md5(open("/tmp/_golden_teapot.ppm", "rb").read()).hexdigest()
    print("raycast -c -w64 -h64 md5:", md5, "(expected a8248b0896d09d1b3b98b67f5c43ce6c)")
