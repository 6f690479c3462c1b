timeout 600 python -m pytest tests/test_gpu_pipeline.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 2>/dev/null | cut -c1-220
