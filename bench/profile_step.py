"""Per-kernel CUDA-event timing of the fused MNIST step (warm L2, as in the real step where the
~60 MB working set stays L2-resident) and the pure graph-replay time with no Python in between.
Writes gpurun_out/profile_step.json."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench as B
import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras
from horovod_distributed_example_b200.ops.gemm import EPI_ATOMIC_ADD, EPI_STORE, EPI_UNPOOL, tc_gemm


def timeit(fn, n=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


def main():
    hvd.init()
    m = B.make_model_ours(hvd, keras, os.environ.get("OPT", "adadelta"))
    m._ensure_built((128, 28, 28, 1))
    e = m._backend
    rng = np.random.default_rng(0)
    x = torch.from_numpy(rng.random((128, 28, 28, 1), dtype=np.float32)).cuda()
    y = torch.from_numpy(rng.integers(0, 10, 128)).cuda()
    for _ in range(5):
        e.train_step(x, y)
    torch.cuda.synchronize()
    lib, s = e.lib, e._s()
    P, G = e._p, e._g
    res = {}
    res["conv1_fwd"] = timeit(lambda: lib.hvd_conv1_fwd(e.x_in.data_ptr(), P("c1", "kernel").data_ptr(), P("c1", "bias").data_ptr(), e.a1.data_ptr(), 128, None, 0, None, 0, e.state.data_ptr(), 0, s))
    res["conv2_fprop"] = timeit(lambda: lib.hvd_conv2_fprop(e.a1.data_ptr(), P("c2", "kernel").data_ptr(), P("c2", "bias").data_ptr(), e.p2.data_ptr(), e.mask2.data_ptr(), e.state.data_ptr(), 128, 1, e.keep1, 0, s))
    for ks in (24, 48, 96, 144):
        res[f"dense1_fwd_ks{ks}"] = timeit(lambda: tc_gemm(e.p2, P("f1", "kernel"), 128, 128, 9216, False, True, 128, ks, EPI_ATOMIC_ADD, out=e.h_pre, stream=s))
    res["head"] = timeit(lambda: e._head(128, True, True, 0))
    for bn in (64, 128):
        res[f"dense1_dx_unpool_bn{bn}"] = timeit(lambda: tc_gemm(e.dh_pre, P("f1", "kernel"), 128, 9216, 128, False, False, bn, epilogue=EPI_UNPOOL, out=e.dy2.view(-1, 64), mask=e.mask2, scale=1.0 / e.keep1, stream=s))
    res["dense1_dw"] = timeit(lambda: tc_gemm(e.p2, e.dh_pre, 9216, 128, 128, True, True, 128, epilogue=EPI_STORE, out=G("f1", "kernel", 0), stream=s))
    res["conv2_dgrad"] = timeit(lambda: lib.hvd_conv2_dgrad(e.dy2.data_ptr(), P("c2", "kernel").data_ptr(), e.a1.data_ptr(), None, 128, 0, s, e.x_in.data_ptr(), G("c1", "kernel", 0).data_ptr(), G("c1", "bias", 0).data_ptr()))
    for ctas in (72, 144):
        res[f"conv2_wgrad_ctas{ctas}"] = timeit(lambda: lib.hvd_conv2_wgrad(e.a1.data_ptr(), e.dy2.data_ptr(), G("c2", "kernel", 0).data_ptr(), G("c2", "bias", 0).data_ptr(), 128, ctas, s))
    if e.fused is not None:
        for i in range(len(e.fused.buckets)):
            res[f"fused_update_bucket{i}"] = timeit(lambda: e.fused.run_bucket(i, 0, 1))
    g = next(iter(e._graphs.values()))
    res["graph_replay_full_step"] = timeit(lambda: g.replay(), n=200)
    res["train_step_python_loop"] = timeit(lambda: e.train_step(x, y), n=200)
    res["sum_of_kernels"] = sum(v for k, v in res.items() if not k.startswith(("graph", "train", "dense1_fwd_ks2", "dense1_fwd_ks4", "dense1_fwd_ks1", "dense1_dx_unpool_bn64", "conv2_wgrad_ctas72")))
    for k, v in res.items():
        print(f"{k:32s} {v:9.2f} us")
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "profile_step.json"), "w"), indent=1)
    hvd.shutdown()


if __name__ == "__main__":
    main()
