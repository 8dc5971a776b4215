#!/bin/bash
# one 8-GPU box session: the scaling curve of bench.py (cfg3 at every N), the driver's torchrun launch shape with all secondary
# workloads, readback-strip variants at 8 GPUs, the multi-GPU tests on all GPUs.  Outputs under gpurun_out/r02_scale/.
out=gpurun_out/r02_scale; mkdir -p $out
python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout 900 > $out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> $out/pytest_multi.log
for n in 1 2 4 8; do
  python bench.py --gpus $n --steps 20 --warmup 5 --extra 0 > $out/bench_n$n.json 2> $out/bench_n$n.err; echo "n=$n rc=$?"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 8 --steps 20 --warmup 5 > $out/bench_torchrun_n8_full.json 2> $out/bench_torchrun_n8_full.err; echo "torchrun rc=$?"
for v in "100000 2" "100000 4" "100000 8"; do set -- $v
  echo "== 8 GPUs, VORTEX_B200_READBACK_MIN_TILES=$1 VORTEX_B200_READBACK_STRIPS=$2" >> $out/strips_n8.txt
  VORTEX_B200_READBACK_MIN_TILES=$1 VORTEX_B200_READBACK_STRIPS=$2 python tools/r02_e2e_probe.py 8 2>&1 | grep -E "full sequence|start\+wait|uploads" >> $out/strips_n8.txt
done
python bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > $out/bench_reference.json 2> $out/bench_reference.err
tail -n 3 $out/pytest_multi.log $out/strips_n8.txt; tail -n 2 $out/*.err
