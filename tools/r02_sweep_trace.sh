#!/bin/bash
o=gpurun_out/r02_sweep_trace_ring_top.txt; : > $o
for lib in - _variants/t_r0_t0 _variants/t_r8_t0 _variants/t_r4_t1 _variants/t_r6_t1 _variants/t_r12_t1 -; do
  l=$lib; [ "$lib" != "-" ] && l=$lib/libvortex-b200.so
  python tools/quick_variant.py $l --workload cfg4_incoherent_10m --no-cpu --extra 0 --steps 20 --warmup 5 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('cfg4', '$lib', 'value', round(d['value'], 1), 'kernel_ms', round(d['roofline']['kernel_ms'], 4), 'parity_green', d['parity']['green'], 'id_mismatch', d['parity']['id_mismatch'], 'non_tie', d['parity']['non_tie'])" >> $o
done
python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale.py -m gpu -q -x --timeout 900 2>&1 | tail -3 >> $o
cat $o
