#!/bin/bash
# kernel-side Mrays/s of cfg4 with the coherence ordering off / on at several Morton widths (gpurun, 1 GPU)
for b in off 5 6 7 8 9; do
  if [ $b = off ]; then export VORTEX_B200_RAY_ORDER=0; else export VORTEX_B200_RAY_ORDER=1 VORTEX_B200_RAY_ORDER_BITS=$b; fi
  timeout 200 python bench.py --workload cfg4_incoherent_10m --no-cpu --steps 10 --warmup 3 2>/dev/null | tail -n 1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('bits=$b', 'value', round(d['value'], 1), 'ms_per_step', round(d['ms_per_step'], 4), 'launches', d['gpu_launches'], 'e2e', round(d['e2e']['value'], 1))"
done
