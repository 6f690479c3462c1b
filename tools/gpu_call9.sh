for bps in 7 8 6; do
echo "== HVD_CONV1_BPS=$bps"
HVD_CONV1_BPS=$bps HVD_TAIL_TRACE=1 timeout 200 python -m horovod_distributed_example_b200.run -np 2 bench/tail_trace.py 300 2>&1 | tail -2 | cut -c1-460
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 300 --warmup 30 --no-baseline 2>/dev/null | cut -c1-250
