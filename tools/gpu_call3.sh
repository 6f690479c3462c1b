set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline.py -x -q > gpurun_out/r2_pipe_test.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pipe_test.log
tail -8 gpurun_out/r2_pipe_test.log
timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline > gpurun_out/r2_bench_N1_pipelined.json 2> gpurun_out/r2_bench_N1_pipelined.err; tail -c 400 gpurun_out/r2_bench_N1_pipelined.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 300 --warmup 30 --no-baseline > gpurun_out/r2_bench_N2_pipelined.json 2> gpurun_out/r2_bench_N2_pipelined.err; tail -c 400 gpurun_out/r2_bench_N2_pipelined.err
cat gpurun_out/r2_bench_N?_pipelined.json | cut -c1-250
timeout 200 python -m horovod_distributed_example_b200.run -np 2 bench/tail_trace.py 300 > gpurun_out/r2_tail_trace_N2.log 2>&1; cat gpurun_out/r2_tail_trace_N2.log | tail -5
timeout 200 python bench/tail_trace.py 300 > gpurun_out/r2_tail_trace_N1.log 2>&1; tail -3 gpurun_out/r2_tail_trace_N1.log
