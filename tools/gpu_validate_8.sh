mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_N8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_N8.log
tail -4 gpurun_out/r2_pytest_gpu_N8.log
run() { n=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $n "$@" 2>/dev/null | tail -1; }
run 8 --steps 300 --warmup 30 > gpurun_out/r2_bench_N8.json; cut -c1-230 gpurun_out/r2_bench_N8.json
run 8 --steps 20 --warmup 5 > gpurun_out/r2_bench_N8_s20.json; cut -c1-230 gpurun_out/r2_bench_N8_s20.json
run 4 --steps 300 --warmup 30 > gpurun_out/r2_bench_N4.json; cut -c1-230 gpurun_out/r2_bench_N4.json
HVD_AG_MC=0 run 8 --steps 300 --warmup 30 --no-baseline > gpurun_out/r2_bench_N8_mc0.json; cut -c1-230 gpurun_out/r2_bench_N8_mc0.json
HVD_TAIL_TRACE=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench/tail_trace.py 300 2>/dev/null | tail -8 > gpurun_out/r2_tail_trace_N8.log; head -2 gpurun_out/r2_tail_trace_N8.log | cut -c1-560
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29537 bench/allreduce_sweep.py --max-mb 256 2>&1 | grep bytes > gpurun_out/r2_sweep_N8.log; cp gpurun_out/allreduce_sweep_N8.json gpurun_out/r2_allreduce_sweep_N8.json; cut -c1-330 gpurun_out/r2_sweep_N8.log
for impl in ours ddp; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29538 bench/resnet50_stress.py --impl $impl --steps 30 --warmup 8 2>/dev/null | tail -1 > gpurun_out/r2_resnet_N8_$impl.json; cut -c1-400 gpurun_out/r2_resnet_N8_$impl.json
done
