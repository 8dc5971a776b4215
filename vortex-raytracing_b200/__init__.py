"""vortex-raytracing_b200 — B200-native backend for the ray-tracing hot path of the Vortex fork
LazyLatte/vortex-raytracing, behind the reference's own vx_* runtime C ABI.

  csrc/       CUDA kernels (sm_100a) + the C-ABI shared library  -> lib/libvortex-b200.so
  host/       C++ host library: scene/BVH/TLAS builders and camera set-up in the reference's buffer layouts
  vx.py       ctypes binding of the vx_* calls (same names as runtime/include/vortex.h)
  tracer.py   the reference Tracer's init/setup/run call sequence over vx_*
  scenes.py   procedural scene generators for the benchmark configs

The directory name contains a hyphen; import it with importlib.import_module("vortex-raytracing_b200").
"""
from . import vx  # noqa: F401
