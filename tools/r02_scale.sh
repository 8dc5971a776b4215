#!/bin/bash
# one 8-GPU box session: the scaling curve of bench.py (cfg3 at every N), the driver's torchrun launch shape, the replication /
# framebuffer variants, and the multi-GPU tests on all GPUs.  Outputs under gpurun_out/r02_scale/.
out=gpurun_out/r02_scale; mkdir -p $out
nvidia-smi topo -m > $out/topo.txt 2>&1
python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout 900 > $out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> $out/pytest_multi.log
for n in 1 2 4 8; do
  python bench.py --gpus $n --steps 20 --warmup 5 --extra 0 > $out/bench_n$n.json 2> $out/bench_n$n.err; echo "n=$n rc=$?"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 8 --steps 20 --warmup 5 > $out/bench_torchrun_n8_full.json 2> $out/bench_torchrun_n8_full.err; echo "torchrun rc=$?"
python bench.py --gpus 8 --steps 20 --warmup 5 --extra 0 --collective nccl > $out/bench_n8_nccl.json 2> $out/bench_n8_nccl.err; echo "nccl rc=$?"
python bench.py --gpus 8 --steps 20 --warmup 5 --extra 0 --fb peer > $out/bench_n8_fbpeer.json 2> $out/bench_n8_fbpeer.err; echo "fbpeer rc=$?"
python bench.py --gpus 8 --steps 10 --warmup 3 --extra 0 --workload cfg4_incoherent_10m > $out/bench_n8_cfg4.json 2> $out/bench_n8_cfg4.err; echo "cfg4 rc=$?"
python bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > $out/bench_reference.json 2> $out/bench_reference.err
tail -3 $out/*.err $out/pytest_multi.log
