#!/bin/bash
# usage: [WLS="..."] [ROUNDS=n] tools/ab_variant.sh <out file> <variant .so> [<variant .so> ...]; kernel-side value of the one-GPU
# workloads, in-tree library ("-") vs each variant, ROUNDS rounds each (A B A B) so that drift shows
out=$1; shift; : > $out
for wl in ${WLS:-cfg1_1080p_primary+shadow_100k cfg2_4k_diffuse4_1m cfg3_8k_primary+shadow_1m cfg4_incoherent_10m}; do
  for round in $(seq 1 ${ROUNDS:-2}); do
    for lib in - "$@"; do
      python tools/quick_variant.py $lib --workload $wl --no-cpu --extra 0 --steps 20 --warmup 5 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$wl', '$lib', 'value', round(d['value'], 1), 'kernel_ms', round(d['roofline']['kernel_ms'], 4), 'fb_off', round(d.get('schedule_feedback_off', {}).get('value', 0), 1), 'e2e', round(d['e2e']['value'], 1), 'parity_green', d['parity']['green'], 'id_mismatch', d['parity']['id_mismatch'], 'non_tie', d['parity']['non_tie'])" >> $out
    done
  done
done
cat $out
