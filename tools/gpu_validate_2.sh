mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_N2d.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_N2d.log
tail -6 gpurun_out/r2_pytest_gpu_N2d.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --impl reference
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/r2_bench_N1_s20.json; cut -c1-200 gpurun_out/r2_bench_N1_s20.json
timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 2>/dev/null | tail -1 > gpurun_out/r2_bench_N1.json; cut -c1-200 gpurun_out/r2_bench_N1.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/r2_bench_N2_s20.json; cut -c1-200 gpurun_out/r2_bench_N2_s20.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 300 --warmup 30 2>/dev/null | tail -1 > gpurun_out/r2_bench_N2.json; cut -c1-200 gpurun_out/r2_bench_N2.json
python bench/fit_throughput.py 2>/dev/null | tail -1
