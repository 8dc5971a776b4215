#!/bin/bash
# ncu evidence of the round (one GPU): launch list of the default bench command + one full capture of the dominant kernel of
# every benched workload.  Reports land in gpurun_out/ncu/; tools/r02_ncu_extract.py turns them into profiles/r02_ncu_*.
o=gpurun_out/ncu; mkdir -p $o
export VORTEX_B200_SCHEDULE=0
python bench.py --steps 2 --warmup 1 --extra 0 --no-cpu > $o/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/r02_launches_cfg3.csv python bench.py --steps 2 --warmup 1 --extra 0 --no-cpu > $o/ncu_bench.log 2>&1
for wl in ${WORKLOADS:-cfg3_8k_primary+shadow_1m cfg2_4k_diffuse4_1m primary_4k_1m cfg1_1080p_primary+shadow_100k cfg4_incoherent_10m rtu_1080p_primary_100k}; do
  python tools/quick_frame.py $wl 5 > $o/plain_$wl.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'k_render_m|k_render_direct|k_rtu_render|k_rtu_direct' -s 3 -c 1 -f -o $o/$wl python tools/quick_frame.py $wl 5 > $o/ncu_$wl.log 2>&1
  echo "$wl rc=$?"
done
ls -la $o
