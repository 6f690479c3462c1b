mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline.py -x -q 2>&1 | tail -3
for fk in 1 0; do
echo "== HVD_RS_FLAG_KERNEL=$fk MC=0"
HVD_RS_FLAG_KERNEL=$fk HVD_TAIL_MC=0 timeout 200 python -m horovod_distributed_example_b200.run -np 2 bench/tail_trace.py 300 2>&1 | tail -2 | cut -c1-420
done
HVD_TAIL_MC=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 300 --warmup 30 --no-baseline 2>/dev/null | cut -c1-250
