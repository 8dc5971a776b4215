"""kernel-only Mrays/s of the benchmark workloads (scratch tool for parameter sweeps): python tools/quick_bench.py [workload ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

names = sys.argv[1:] or ["cfg1_1080p_primary+shadow_100k", "primary_4k_1m", "cfg2_4k_diffuse4_1m"]
out = []
for n in names:
    r = bench.quick_run(n, 0, steps=5, kernel=int(os.environ.get("VORTEX_B200_KERNEL", "0")))
    out.append(f"{n.split('_')[0]}={r['mrays_per_s']:.0f}")
print(os.environ.get("TAG", ""), " ".join(out), flush=True)
