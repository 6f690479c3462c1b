"""``keras.utils`` pieces the reference uses: ``to_categorical`` (``mnist_keras.py:68-69``); plus the lazy
scalar that lets per-batch logs reach callbacks without a device sync."""
import numpy as np


class LazyScalar:
    """A per-batch metric that is still on its way from the device.

    ``fit`` hands these to callbacks in ``logs`` when the backend publishes step results asynchronously.
    ``float(v)`` (also ``format`` / ``repr`` / ``item()``) waits for that one step; ``v.ready()`` tells a
    consumer whether that would block, so writers such as ``TensorBoard(update_freq='batch')`` emit a batch's
    scalars one batch late instead of stalling the launch of the next step."""

    __slots__ = ("_resolve", "_ready", "_value")

    def __init__(self, resolve, ready):
        self._resolve, self._ready, self._value = resolve, ready, None

    def ready(self) -> bool:
        return self._value is not None or bool(self._ready())

    def __float__(self) -> float:
        if self._value is None:
            self._value = float(self._resolve())
            self._resolve = self._ready = None
        return self._value

    def item(self) -> float:
        return float(self)

    def __format__(self, spec) -> str:
        return format(float(self), spec)

    def __repr__(self) -> str:
        return repr(float(self))


def is_ready(v) -> bool:
    """False only for a LazyScalar whose step has not finished yet."""
    r = getattr(v, "ready", None)
    return True if r is None else bool(r())


def to_categorical(y, num_classes=None, dtype="float32"):
    y = np.asarray(y, dtype="int64").reshape(-1)
    if num_classes is None:
        num_classes = int(y.max()) + 1
    out = np.zeros((y.shape[0], num_classes), dtype=dtype)
    out[np.arange(y.shape[0]), y] = 1
    return out
