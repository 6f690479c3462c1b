mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_N8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_N8.log
tail -6 gpurun_out/r2_pytest_gpu_N8.log
