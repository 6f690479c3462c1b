python bench.py --impl reference
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/final_bench_N2_s20.json; cut -c1-230 gpurun_out/final_bench_N2_s20.json
