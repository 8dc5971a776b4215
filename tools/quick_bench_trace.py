"""one-GPU run of an incoherent-ray workload without the CPU leg (ncu target): python tools/quick_bench_trace.py [workload]"""
import os
import subprocess
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg4_incoherent_10m"
sys.argv = ["bench.py", "--workload", wl, "--steps", "3", "--warmup", "3", "--no-cpu"]
sys.path.insert(0, root)
import bench  # noqa: E402

bench.main()
