mkdir -p gpurun_out
HVD_NO_GRAPH=1 python bench/eager_step.py 6 > gpurun_out/plain_eager.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'conv|tc_gemm|head|hvd_' -s 36 -c 9 -f -o gpurun_out/prof_step_r2 python bench/eager_step.py 6 > gpurun_out/ncu_eager.log 2>&1
tail -3 gpurun_out/plain_eager.log; tail -5 gpurun_out/ncu_eager.log
for opt in "--optimizer adam --lr-warmup" "--optimizer adadelta --lr-warmup" "--optimizer adadelta"; do
timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline $opt 2>/dev/null | cut -c1-200
done
python bench/profile_step.py 2>&1 | tail -25
