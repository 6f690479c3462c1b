mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu_N2d.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_gpu_N2d.log
tail -4 gpurun_out/r2_pytest_gpu_N2d.log
