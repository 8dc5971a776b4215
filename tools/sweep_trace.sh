#!/bin/bash
# sweep of the state machine's scheduling parameters for the incoherent-ray workload (scratch tool)
# usage: RF="16 24" QI="12 16" CH="64 128" tools/sweep_trace.sh [workload]
for rf in ${RF:-8 12 16 24}; do for qi in ${QI:-4 8 12}; do
  [ $qi -ge $rf ] && continue
  for ch in ${CH:-32 64}; do
    v=$(VORTEX_B200_TRACE_QUORUM_REFILL=$rf VORTEX_B200_TRACE_QUORUM_INNER=$qi VORTEX_B200_TRACE_CHUNK=$ch python bench.py --workload ${1:-incoherent_1m} --steps 5 --warmup 3 --no-cpu 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().splitlines()[-1]); print(round(d['value']))")
    echo "refill=$rf inner=$qi chunk=$ch mrays=$v"
  done
done; done
