"""A few EAGER (HVD_NO_GRAPH=1) training steps of the fused engine on one GPU - the target of the ncu capture
(kernels inside a replayed CUDA graph are awkward to select by launch index):

    HVD_NO_GRAPH=1 python bench/eager_step.py 6
    ncu --set full --clock-control none --import-source on -k regex:'conv|tc_gemm|head|hvd_' -s 36 -c 9 \
        -o gpurun_out/prof_step_r2 python bench/eager_step.py 6"""
import os
import sys

os.environ.setdefault("HVD_NO_GRAPH", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np                                              # noqa: E402
import torch                                                    # noqa: E402

import bench as B                                               # noqa: E402
import horovod_distributed_example_b200 as hvd                  # noqa: E402
from horovod_distributed_example_b200 import keras              # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    hvd.init()
    m = B.make_model_ours(hvd, keras, os.environ.get("OPT", "adadelta"))
    rng = np.random.default_rng(0)
    x = rng.random((128, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, 128)
    for _ in range(steps):
        out = m.train_on_batch(x, y)
    torch.cuda.synchronize()
    print("eager steps done, loss", out[0], "launches/step", m._backend.launches_per_step)
    hvd.shutdown()


if __name__ == "__main__":
    main()
