mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_collectives.py -x -q 2>&1 | tail -4
HVD_TAIL_TRACE=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench/tail_trace.py 300 2>&1 | tail -2 | cut -c1-560
for cfg in "HVD_AG_WHERE=conv1" "HVD_AG_WHERE=start" "HVD_AG_WHERE=conv1 HVD_AG_CTAS=64" "HVD_AG_WHERE=conv1 HVD_AG_MC=1"; do
echo "== $cfg"
env $cfg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 300 --warmup 30 --no-baseline 2>/dev/null | cut -c1-250
done
timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline 2>/dev/null | cut -c1-250
