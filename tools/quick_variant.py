"""A/B of a kernel build variant: python tools/quick_variant.py <path to an alternative libvortex-b200.so | -> <bench.py arguments...>
("-" = the in-tree library).  Variants are built with `make -C vortex-raytracing_b200/csrc OUT=<dir> EXTRA=-D...`."""
import os
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
alt = sys.argv[1]
sys.argv = ["bench.py"] + sys.argv[2:]
sys.path.insert(0, root)
import bench  # noqa: E402

if alt != "-":
    bench.pkg("vx").LIB_PATH = os.path.abspath(alt)
bench.main()
