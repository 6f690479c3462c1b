mkdir -p gpurun_out
HVD_TEST_BODY=torchmodel_bn_vs_ddp timeout 300 python -m horovod_distributed_example_b200.run -np 2 tests/_gpu_workers.py 2>&1 | tail -25
for impl in ours ddp; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29538 bench/resnet50_stress.py --impl $impl --steps 30 --warmup 8 2>/dev/null | tail -1 > gpurun_out/r2_resnet_N2_$impl.json; cut -c1-900 gpurun_out/r2_resnet_N2_$impl.json
done
