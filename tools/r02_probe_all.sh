#!/bin/bash
out=gpurun_out/r02_e2e_probe3.txt; : > $out
python -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout 900 2>&1 | tail -3 >> $out
for n in 2 8; do python tools/r02_e2e_probe.py $n >> $out 2>&1; echo >> $out; done
echo "VORTEX_B200_GROUP_THREADS=0" >> $out; VORTEX_B200_GROUP_THREADS=0 python tools/r02_e2e_probe.py 8 >> $out 2>&1
python bench.py --gpus 8 --steps 20 --warmup 5 --extra 0 > gpurun_out/r02_n8_v3.json 2> gpurun_out/r02_n8_v3.err; echo "bench8 rc=$?" >> $out
cat $out
