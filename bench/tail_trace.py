"""Where the pipelined step's tail kernel spends its time, from its own %globaltimer stamps.

    HVD_TAIL_TRACE=1 python -m horovod_distributed_example_b200.run -np N bench/tail_trace.py [steps]

Prints, per rank, the mean over the last 48 steps of (microseconds, relative to the kernel's entry):
LL part: packets pushed / every peer's packets arrived / done; shard part: tiles there / peers entered / published;
and the step period.  See csrc/comm/pipeline_tail.cu for the stamp positions."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as B                                              # noqa: E402
import horovod_distributed_example_b200 as hvd                 # noqa: E402
from horovod_distributed_example_b200 import keras             # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    os.environ["HVD_TAIL_TRACE"] = "1"
    hvd.init()
    model = B.make_model_ours(hvd, keras, os.environ.get("OPT", "adadelta"))
    model._ensure_built((128, 28, 28, 1))
    hvd.broadcast_global_variables(0, model)
    be = model._backend
    rng = np.random.default_rng(hvd.rank())
    x = torch.from_numpy(rng.random((16, 128, 28, 28, 1), dtype=np.float32)).pin_memory().numpy()
    y = torch.from_numpy(rng.integers(0, 10, (16, 128)).astype(np.int64)).pin_memory().numpy()
    for i in range(steps):
        be.train_host_lazy(x[i % 16], y[i % 16])
    torch.cuda.synchronize()
    hvd.barrier()
    tr = be.pipe.read_trace().numpy().astype(np.int64)
    order = np.argsort(tr[:, 0])
    tr = tr[order][-49:-1]                  # (the newest row has no all-gather / next-step stamps yet)
    rel = lambda k, b: float(np.median(tr[:, k] - tr[:, b])) / 1e3
    out = {"rank": hvd.rank(), "n": hvd.size(),
           "ll_pushed": rel(1, 0), "ll_all_arrived": rel(2, 0), "ll_done": rel(3, 0),
           "shard_entry_vs_ll_entry": rel(4, 0), "shard_tiles_there": rel(5, 4), "shard_done": rel(7, 4),
           "ag_start_after_tail_entry": rel(10, 0), "ag_flags_after_tail_entry": rel(11, 0),
           "d1fwd_wait_begin_after_tail_entry": rel(12, 0), "d1fwd_wait_us": rel(13, 12), "step_period": float(np.median(np.diff(tr[:, 0]))) / 1e3,
           "entry_abs_last": int(tr[-1, 0])}
    allo = hvd.allgather_object(out)
    if hvd.rank() == 0:
        for o in allo:
            print(json.dumps(o))
    hvd.shutdown()


if __name__ == "__main__":
    main()
