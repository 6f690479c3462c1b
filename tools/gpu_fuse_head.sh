HVD_FUSE_HEAD=1 timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -k "engine" 2>&1 | tail -4
for v in 0 1; do echo "== HVD_FUSE_HEAD=$v"; HVD_FUSE_HEAD=$v timeout 200 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline --no-e2e 2>/dev/null | cut -c60-200; done
