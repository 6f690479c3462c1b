#!/bin/bash
# Round-end style evidence on N GPUs: GPU tests, headline bench for ours and the stock baseline, ResNet-50 stress.
N=${1:-8}
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu_N$N.log; tail -4 gpurun_out/pytest_gpu_N$N.log
for impl in ours torch_nccl; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$N bench.py --gpus $N --steps 300 --warmup 30 --impl $impl > gpurun_out/bench_${impl}_N$N.json 2> gpurun_out/bench_${impl}_N$N.err
  tail -1 gpurun_out/bench_${impl}_N$N.json | cut -c1-250; grep -E "Error" gpurun_out/bench_${impl}_N$N.err | tail -3
done
for impl in ours ddp; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$N bench/resnet50_stress.py --impl $impl --steps 15 --warmup 5 2>&1 | grep -E "^\{|Error" | tail -2 | tee -a gpurun_out/resnet_N$N.jsonl
done
