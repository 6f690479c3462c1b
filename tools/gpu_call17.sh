mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_collectives.py -x -q 2>&1 | tail -4
python bench/tf32_peak.py > gpurun_out/tf32_peak.json 2>/dev/null; cat gpurun_out/tf32_peak.json
for opt in "--optimizer adadelta" "--optimizer adam" "--optimizer adam --lr-warmup" "--optimizer adadelta --lr-warmup"; do
tag=$(echo $opt | tr -d ' -' )
timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline $opt 2>/dev/null > gpurun_out/r2_bench_N1_$tag.json; cut -c1-200 gpurun_out/r2_bench_N1_$tag.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 300 --warmup 30 --no-baseline $opt 2>/dev/null > gpurun_out/r2_bench_N2_$tag.json; cut -c1-200 gpurun_out/r2_bench_N2_$tag.json
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29537 bench/allreduce_sweep.py --max-mb 256 2>&1 | grep bytes | cut -c1-400
cp gpurun_out/allreduce_sweep_N2.json gpurun_out/r2_allreduce_sweep_N2.json
