"""Throughput of the user-facing training loop: Model.fit on 60,000 MNIST-shaped samples, batch 128, with
the callbacks the reference scripts use (TensorBoard with per-batch scalars on rank 0).  1 GPU."""
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench as B
import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras


def main():
    hvd.init()
    rng = np.random.default_rng(0)
    x = rng.random((60000, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, 60000).astype(np.int64)
    out = {}
    for name, make_cbs in (("no_callbacks", lambda d: []),
                           ("tensorboard_batch", lambda d: [keras.callbacks.TensorBoard(log_dir=d, update_freq="batch")])):
        model = B.make_model_ours(hvd, keras, "adadelta")
        with tempfile.TemporaryDirectory() as d:
            model.fit(x[:4096 + 96], y[:4096 + 96], batch_size=128, epochs=2, verbose=0, callbacks=make_cbs(d))   # warm-up: captures the
            # graphs of the full batch and of the 96-sample tail (60000 = 468 * 128 + 96), both step parities
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            model.fit(x, y, batch_size=128, epochs=2, verbose=0, callbacks=make_cbs(d))
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        out[name] = {"images_per_sec": round(2 * 60000 / dt), "us_per_step": round(dt / (2 * 469) * 1e6, 1)}
    print(json.dumps(out))
    hvd.shutdown()


if __name__ == "__main__":
    main()
