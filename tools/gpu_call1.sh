set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_topo.txt
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_N2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_N2.log
tail -5 gpurun_out/r2_pytest_gpu_N2.log
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_N1_s20.json 2> gpurun_out/r2_bench_N1_s20.err; tail -c 600 gpurun_out/r2_bench_N1_s20.err
python bench.py --gpus 1 --steps 300 --warmup 30 > gpurun_out/r2_bench_N1.json 2> gpurun_out/r2_bench_N1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_N2_s20.json 2> gpurun_out/r2_bench_N2_s20.err; tail -c 600 gpurun_out/r2_bench_N2_s20.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 300 --warmup 30 > gpurun_out/r2_bench_N2.json 2> gpurun_out/r2_bench_N2.err
cat gpurun_out/r2_bench_N1_s20.json gpurun_out/r2_bench_N1.json gpurun_out/r2_bench_N2_s20.json gpurun_out/r2_bench_N2.json | cut -c1-700
