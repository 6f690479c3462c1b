timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q 2>&1 | tail -3
timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline 2>/dev/null | cut -c1-200
timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline 2>/dev/null | cut -c1-200
