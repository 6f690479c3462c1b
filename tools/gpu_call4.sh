mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline.py -x -q 2>&1 | tail -8
for mc in 1 0; do
echo "== HVD_TAIL_MC=$mc"
HVD_TAIL_MC=$mc timeout 200 python -m horovod_distributed_example_b200.run -np 2 bench/tail_trace.py 300 2>&1 | tail -2
done
for c in 148 296; do
echo "== HVD_TAIL_CTAS=$c MC=0"
HVD_TAIL_MC=0 HVD_TAIL_CTAS=$c timeout 200 python -m horovod_distributed_example_b200.run -np 2 bench/tail_trace.py 300 2>&1 | tail -1
done
