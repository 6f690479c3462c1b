"""Where does a bench step's time go?  Device time (CUDA events) and host enqueue time (perf_counter, no sync
inside the loop) for: bare graph replay, _run_step (parity alternation, hyper push, watchdog), train_step
(+ in-stream D2D staging of the batch) and the public async host path.  1 GPU; prints one JSON object."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench as B
import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras


def timed(fn, n):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for i in range(n):
        fn(i)
    t1 = time.perf_counter()
    e1.record()
    torch.cuda.synchronize()
    return {"device_us": e0.elapsed_time(e1) * 1e3 / n, "host_us": (t1 - t0) * 1e6 / n}


def main():
    hvd.init()
    dev = hvd.device()
    model = B.make_model_ours(hvd, keras, "adadelta")
    model._ensure_built((128, 28, 28, 1))
    be = model._backend
    rng = np.random.default_rng(0)
    pool = 416
    xs = torch.from_numpy(rng.random((pool, 128, 784), dtype=np.float32)).pin_memory()
    ys = torch.from_numpy(rng.integers(0, 10, (pool, 128)).astype(np.int64)).pin_memory()
    xd, yd = xs.to(dev), ys.to(dev)
    for i in range(10):
        be.train_step(xd[i], yd[i])
    n = int(os.environ.get("N", 400))
    out = {}
    g0 = be._graphs[(128, 0)]
    out["graph_replay_same_parity"] = timed(lambda i: g0.replay(), n)
    gs = [be._graphs[(128, 0)], be._graphs[(128, 1)]]
    out["graph_replay_alternating"] = timed(lambda i: gs[i & 1].replay(), n)
    out["run_step"] = timed(lambda i: be._run_step(128), n)
    out["train_step_dev_pool"] = timed(lambda i: be.train_step(xd[i % pool], yd[i % pool]), n)
    xh, yh = xs.numpy(), ys.numpy()
    out["train_host_lazy"] = timed(lambda i: be.train_host_lazy(xh[i % pool], yh[i % pool]), n)
    print(json.dumps(out, indent=1))
    hvd.shutdown()


if __name__ == "__main__":
    main()
