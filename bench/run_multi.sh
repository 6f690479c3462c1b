#!/bin/bash
# Multi-GPU evidence run: GPU tests, allreduce sweep, scaling bench for ours and the stock baseline.
# usage: bench/run_multi.sh <ngpus>
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu_N$N.log; cat gpurun_out/pytest_gpu_N$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench/allreduce_sweep.py --max-mb 1024 > gpurun_out/sweep_N$N.log 2>&1; grep -E "^\{" gpurun_out/sweep_N$N.log | cut -c1-330 | tail -12; grep -iE "error|Traceback" gpurun_out/sweep_N$N.log | head -5
for n in $N 4; do
  if [ $n -le $N ]; then
    for impl in ours torch_nccl; do
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n bench.py --gpus $n --steps 300 --warmup 30 --impl $impl > gpurun_out/bench_${impl}_N$n.json 2> gpurun_out/bench_${impl}_N$n.err
      tail -1 gpurun_out/bench_${impl}_N$n.json | cut -c1-260; grep -E "Error" gpurun_out/bench_${impl}_N$n.err | tail -3
    done
  fi
done
