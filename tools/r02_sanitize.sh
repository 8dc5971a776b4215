#!/bin/bash
# compute-sanitizer (memcheck, then racecheck on the shared-memory kernels) over small cases of every kernel family
o=gpurun_out/r02_sanitizer.txt; : > $o
run() { echo "== $*" >> $o; timeout 900 "$@" 2>&1 | grep -E "ERROR SUMMARY|passed|failed|Invalid|RACECHECK SUMMARY|hazard|error" | tail -8 >> $o; }
run compute-sanitizer --tool memcheck --print-limit 5 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "cornell64_cfg0 or ragged_37x19 or one_pixel or multi_shadow_diffuse or deep_tree or headroom"
run compute-sanitizer --tool memcheck --print-limit 5 python -m pytest tests/test_gpu_rtu.py -m gpu -q -x -k "cornell64 or ragged or one_pixel"
run compute-sanitizer --tool memcheck --print-limit 5 python -m pytest tests/test_gpu_build.py -m gpu -q -x -k "cornell"
run compute-sanitizer --tool racecheck --print-limit 5 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "cornell64_cfg0 or multi_shadow_diffuse"
cat $o
