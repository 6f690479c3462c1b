set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline.py -x -q > gpurun_out/r2_pipe_test.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pipe_test.log
tail -30 gpurun_out/r2_pipe_test.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_N2b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_N2b.log
tail -15 gpurun_out/r2_pytest_gpu_N2b.log
for mode in pipelined tail; do
HVD_BUCKET_MODE=$mode timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline > gpurun_out/r2_bench_N1_$mode.json 2> gpurun_out/r2_bench_N1_$mode.err; tail -c 400 gpurun_out/r2_bench_N1_$mode.err
HVD_BUCKET_MODE=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 300 --warmup 30 --no-baseline > gpurun_out/r2_bench_N2_$mode.json 2> gpurun_out/r2_bench_N2_$mode.err; tail -c 400 gpurun_out/r2_bench_N2_$mode.err
done
cat gpurun_out/r2_bench_N?_pipelined.json gpurun_out/r2_bench_N?_tail.json | cut -c1-330
