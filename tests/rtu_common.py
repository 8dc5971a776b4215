"""Shared helpers for the RT-unit (layout B) tests: scenes, oracle selection, tracer construction."""
import importlib

import numpy as np

from rt_common import scenes

rtx_tracer = importlib.import_module("vortex-raytracing_b200.rtx_tracer")


def rtu():
    return importlib.import_module("oracle.pyoracle_rtu")


def translate_scale(t, s):
    m = np.eye(4, dtype=np.float32)
    m[0, 0] = m[1, 1] = m[2, 2] = s
    m[0, 3], m[1, 3], m[2, 3] = t
    return m


def b_mesh(mesh, diffuse=None, transform=None, reflectivity=None, two_materials=False):
    """layout-A procedural mesh -> layout-B mesh dict (materials instead of one per-instance texture)"""
    R = rtu()
    m = R.from_layout_a(mesh, diffuse)
    if transform is not None:
        m["transform"] = transform
    if reflectivity is not None:
        m["reflectivity"] = reflectivity
    if two_materials:   # second, untextured material on every other triangle
        m["materials"] = [m["materials"][0], {"diffuse": (0.9, 0.3, 0.2)}]
        m["texid"] = (np.arange(m["tris"].shape[0]) % 2).astype(np.uint32)
    return m


# name -> (mesh factory, width, height, max_depth, camera)
SCENES_B = {
    "cornell64": (lambda: [b_mesh(scenes.cornell_box(), two_materials=True)], 64, 64, 1, 1),          # flat walls: zero-extent quantization
    "hf32": (lambda: [b_mesh(scenes.heightfield(32))], 96, 72, 1, 1),
    "multi3_mirror": (lambda: [b_mesh(scenes.heightfield(16)),
                               b_mesh(scenes.icosphere((0, 0.5, 0), 0.4, 2), diffuse=(0.2, 0.6, 0.9), reflectivity=0.5),
                               b_mesh(scenes.cornell_box(), reflectivity=0.3, transform=translate_scale((2.5, 0, 0), 1.0))],
                      96, 72, 3, 1),
    "fixed_camera": (lambda: [b_mesh(scenes.icosphere((0, 0, 0), 1.0, 3), transform=translate_scale((40.0, 100.0, 0.0), 12.0)),
                              b_mesh(scenes.heightfield(24), transform=translate_scale((60.0, 80.0, 0.0), 40.0))],
                     80, 48, 1, 0),                                                                    # kernel.cpp:28-39 camera
    "ragged_37x19": (lambda: [b_mesh(scenes.heightfield(8))], 37, 19, 1, 1),
    "one_pixel": (lambda: [b_mesh(scenes.cornell_box())], 1, 1, 1, 1),
}


def tracer_b(bufs, karg, camera, cuda_device=None):
    w, h = int(karg["dst_width"][0]), int(karg["dst_height"][0])
    tr = rtx_tracer.RtxTracer(w, h, int(karg["samples_per_pixel"][0]), int(karg["max_depth"][0]), cuda_device=cuda_device)
    tr.init(bufs)
    tr.setup({k: karg[k][0] for k in karg.dtype.names}, camera=camera)
    return tr
