mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_N8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_N8.log
tail -4 gpurun_out/r2_pytest_gpu_N8.log
for mc in 1 0; do
echo "== N=8 HVD_TAIL_MC=$mc"
HVD_TAIL_MC=$mc timeout 200 python -m horovod_distributed_example_b200.run -np 8 bench/tail_trace.py 300 2>&1 | tail -8 | cut -c1-400 > gpurun_out/r2_tail_trace_N8_mc$mc.log; head -2 gpurun_out/r2_tail_trace_N8_mc$mc.log
HVD_TAIL_MC=$mc timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 8 --steps 300 --warmup 30 --no-baseline 2>/dev/null > gpurun_out/r2_bench_N8_mc$mc.json; cut -c1-250 gpurun_out/r2_bench_N8_mc$mc.json
done
for mc in 1 0; do
echo "== N=4 HVD_TAIL_MC=$mc"
HVD_TAIL_MC=$mc timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus 4 --steps 300 --warmup 30 --no-baseline 2>/dev/null > gpurun_out/r2_bench_N4_mc$mc.json; cut -c1-250 gpurun_out/r2_bench_N4_mc$mc.json
done
HVD_BUCKET_MODE=tail timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29536 bench.py --gpus 8 --steps 300 --warmup 30 --no-baseline 2>/dev/null > gpurun_out/r2_bench_N8_legacy.json; cut -c1-250 gpurun_out/r2_bench_N8_legacy.json
