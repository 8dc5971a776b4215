#!/bin/bash
# usage: tools/ab_variant.sh <variant .so> <out file>; kernel-side value of the four one-GPU workloads, in-tree library vs variant
alt=$1; out=$2; : > $out
for wl in cfg1_1080p_primary+shadow_100k cfg2_4k_diffuse4_1m cfg3_8k_primary+shadow_1m cfg4_incoherent_10m; do
  for lib in - $alt; do
    python tools/quick_variant.py $lib --workload $wl --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl', '$lib', 'value', round(d['value'], 1), 'kernel_ms', round(d['roofline']['kernel_ms'], 4))" >> $out
  done
done
cat $out
