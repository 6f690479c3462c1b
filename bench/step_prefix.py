"""In-situ critical-path profile of the captured training step: re-capture the graph with only a PREFIX of
the step's kernels (in dependency order) and time warm back-to-back replays; the difference between
consecutive prefixes is what that kernel adds to the step where it actually runs (warm L2, PDL overlap,
concurrent branches) - which per-kernel ncu timings (cold, serialised) cannot show.  1 GPU."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench as B
import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras

ORDER = ["conv1_fwd", "fprop", "d1fwd", "head", "dx", "dw", "b0", "dgrad", "wgrad", "b1"]


def main():
    hvd.init()
    dev = hvd.device()
    model = B.make_model_ours(hvd, keras, os.environ.get("OPT", "adadelta"))
    model._ensure_built((128, 28, 28, 1))
    be = model._backend
    rng = np.random.default_rng(0)
    x = torch.from_numpy(rng.random((128, 784), dtype=np.float32)).to(dev)
    y = torch.from_numpy(rng.integers(0, 10, (128,)).astype(np.int64)).to(dev)
    for _ in range(5):
        be.train_step(x, y)
    n = int(os.environ.get("N", 300))

    def measure(keep):
        be._SKIP = set(ORDER) - set(keep)
        be._graphs.clear()
        torch.cuda.synchronize()
        g = be._capture(128, 0)
        for _ in range(10):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / n

    rows = []
    prev = 0.0
    for k in range(1, len(ORDER) + 1):
        t = measure(ORDER[:k])
        rows.append({"prefix_ends_with": ORDER[k - 1], "replay_us": round(t, 2), "added_us": round(t - prev, 2)})
        prev = t
    extra = {
        "full_without_wgrad": round(measure([o for o in ORDER if o != "wgrad"]), 2),
        "full_without_dgrad": round(measure([o for o in ORDER if o != "dgrad"]), 2),
        "full_without_b0": round(measure([o for o in ORDER if o != "b0"]), 2),
        "full_without_dw_b0": round(measure([o for o in ORDER if o not in ("b0", "dw")]), 2),
        "forward_only": round(measure(ORDER[:4]), 2),
    }
    print(json.dumps({"prefixes": rows, "variants": extra}, indent=1))
    hvd.shutdown()


if __name__ == "__main__":
    main()
