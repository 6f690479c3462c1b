mkdir -p gpurun_out
run() { n=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $n "$@" 2>/dev/null | tail -1; }
run 8 --steps 300 --warmup 30 > gpurun_out/r2_bench_N8.json; cut -c1-200 gpurun_out/r2_bench_N8.json
run 8 --steps 20 --warmup 5 > gpurun_out/r2_bench_N8_s20.json; cut -c1-200 gpurun_out/r2_bench_N8_s20.json
run 4 --steps 300 --warmup 30 > gpurun_out/r2_bench_N4.json; cut -c1-200 gpurun_out/r2_bench_N4.json
