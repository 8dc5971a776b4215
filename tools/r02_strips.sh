#!/bin/bash
o=gpurun_out/r02_sweep_readback_strips.txt; : > $o
for wl in cfg3_8k_primary+shadow_1m cfg1_1080p_primary+shadow_100k; do
  for s in 1 2 4 6 8; do
    echo "== $wl VORTEX_B200_READBACK_STRIPS=$s" >> $o
    VORTEX_B200_READBACK_STRIPS=$s python tools/r02_e2e_probe.py ${N:-1} $wl 2>&1 | grep -E "full sequence|start\+wait" >> $o
  done
done
cat $o
