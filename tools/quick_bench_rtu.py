"""kernel-only Mrays/s of the layout-B (RT-unit) workload (scratch tool / ncu target): python tools/quick_bench_rtu.py [n w h]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

a = [int(x) for x in sys.argv[1:4]] or [224, 1920, 1080]
r = bench.quick_run_rtu(0, n=a[0], w=a[1], h=a[2], with_cpu=False)
print(os.environ.get("TAG", ""), json.dumps(r), flush=True)
