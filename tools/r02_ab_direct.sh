#!/bin/bash
o=gpurun_out/r02_sweep_direct_a.txt; : > $o
VORTEX_B200_DIRECT=1 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale.py -m gpu -q --timeout 900 2>&1 | tail -4 >> $o
for wl in cfg1_1080p_primary+shadow_100k cfg2_4k_diffuse4_1m cfg3_8k_primary+shadow_1m primary_4k_1m; do
  for d in 0 1 0 1; do
    VORTEX_B200_DIRECT=$d python bench.py --workload $wl --no-cpu --extra 0 --steps 20 --warmup 5 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl', 'direct=$d', 'value', round(d['value'], 1), 'kernel_ms', round(d['roofline']['kernel_ms'], 4), 'fb_off', round(d.get('schedule_feedback_off', {}).get('value', 0), 1), 'e2e', round(d['e2e']['value'], 1), 'parity_green', d['parity']['green'], 'id_mismatch', d['parity']['id_mismatch'], 'non_tie', d['parity']['non_tie'])" >> $o
  done
done
cat $o
