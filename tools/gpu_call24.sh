for cfg in "HVD_D1_KSPLIT=16" "HVD_D1_KSPLIT=48" "HVD_D1_KSPLIT=72" "HVD_D1_KSPLIT=144" "HVD_TAIL_CTAS=222" "HVD_LL_CTAS=8" "HVD_DGRAD_CTAS=112 HVD_WGRAD_CTAS=36" "HVD_DGRAD_CTAS=108 HVD_WGRAD_CTAS=40"; do
echo "== $cfg"
env $cfg timeout 300 python bench.py --gpus 1 --steps 300 --warmup 30 --no-baseline --no-e2e 2>/dev/null | cut -c60-190
done
