"""A small ``tf.data``-like pipeline and array batching.

Reference use (``tensorflow2_keras_mnist.py:37-41``)::

    dataset = tf.data.Dataset.from_tensor_slices((images, labels))
    dataset = dataset.repeat().shuffle(10000).batch(128)

Only the operators the reference touches are provided (plus ``take``/``map``).  The
pipeline works on *indices* - a repeat/shuffle-buffer/batch stream of row indices - and
gathers rows from the source arrays once per batch, so a batch costs one fancy-index
per array.  Batches are produced into pinned host memory when CUDA is available so the
trainer's per-step H2D copy is asynchronous.
"""
from __future__ import annotations

import os
from typing import Callable, Iterator, Optional, Sequence, Tuple

import numpy as np


class Dataset:
    def __init__(self, arrays: Tuple[np.ndarray, ...]):
        n = len(arrays[0])
        for a in arrays:
            if len(a) != n:
                raise ValueError("all arrays must share the first dimension")
        self._arrays = tuple(np.asarray(a) for a in arrays)
        self._n = n
        self._ops = []          # list of (op, arg)
        self._seed = None

    @staticmethod
    def from_tensor_slices(tensors) -> "Dataset":
        if isinstance(tensors, (tuple, list)):
            return Dataset(tuple(_to_numpy(t) for t in tensors))
        return Dataset((_to_numpy(tensors),))

    def _derive(self, op, arg) -> "Dataset":
        d = Dataset.__new__(Dataset)
        d._arrays, d._n, d._seed = self._arrays, self._n, self._seed
        d._ops = self._ops + [(op, arg)]
        return d

    def repeat(self, count: Optional[int] = None) -> "Dataset":
        return self._derive("repeat", count)

    def shuffle(self, buffer_size: int, seed: Optional[int] = None) -> "Dataset":
        d = self._derive("shuffle", int(buffer_size))
        if seed is not None:
            d._seed = seed
        return d

    def batch(self, batch_size: int, drop_remainder: bool = False) -> "Dataset":
        return self._derive("batch", (int(batch_size), drop_remainder))

    def take(self, count: int) -> "Dataset":
        return self._derive("take", int(count))

    def map(self, fn: Callable) -> "Dataset":
        return self._derive("map", fn)

    @property
    def cardinality(self) -> int:
        """Number of elements produced, or -1 if infinite."""
        n = self._n
        for op, arg in self._ops:
            if op == "repeat":
                if arg is None:
                    return -1
                n *= arg
            elif op == "batch":
                bs, drop = arg
                if n >= 0:
                    n = n // bs if drop else -(-n // bs)
            elif op == "take":
                n = arg if n < 0 else min(n, arg)
        return n

    # -- index stream ----------------------------------------------------------
    def _index_stream(self, rng) -> Iterator:
        def base():
            for i in range(self._n):
                yield i
        stream = base()
        batched = False
        maps = []
        for op, arg in self._ops:
            if op == "repeat":
                stream = _repeat(self, stream, arg, self._ops, rng)
            elif op == "shuffle":
                stream = _shuffle(stream, arg, rng)
            elif op == "batch":
                stream = _batch(stream, *arg)
                batched = True
            elif op == "take":
                stream = _take(stream, arg)
            elif op == "map":
                maps.append(arg)
        return stream, batched, maps

    def __iter__(self):
        rng = np.random.default_rng(self._seed)
        stream, batched, maps = self._index_stream(rng)
        for idx in stream:
            if batched:
                idx = np.asarray(idx, dtype=np.int64)
            out = tuple(a[idx] for a in self._arrays)
            for fn in maps:
                out = fn(*out)
                if not isinstance(out, tuple):
                    out = (out,)
            yield out if len(out) > 1 else out[0]


def _repeat(ds, stream, count, ops, rng):
    # ``repeat`` restarts everything upstream of it; upstream here is always the plain
    # range (the reference calls repeat() first), so restart the base range.
    first = True
    n = 0
    while count is None or n < count:
        if first:
            yield from stream
            first = False
        else:
            yield from range(ds._n)
        n += 1


def _shuffle(stream, buffer_size, rng):
    buf = []
    for i in stream:
        if len(buf) < buffer_size:
            buf.append(i)
            continue
        j = int(rng.integers(0, buffer_size))
        out, buf[j] = buf[j], i
        yield out
    rng.shuffle(buf)
    yield from buf


def _batch(stream, bs, drop):
    cur = []
    for i in stream:
        cur.append(i)
        if len(cur) == bs:
            yield cur
            cur = []
    if cur and not drop:
        yield cur


def _take(stream, n):
    for k, i in enumerate(stream):
        if k >= n:
            return
        yield i


def _to_numpy(t):
    try:
        import torch
        if isinstance(t, torch.Tensor):
            return t.detach().cpu().numpy()
    except ImportError:  # pragma: no cover
        pass
    return np.asarray(t)


class ArrayBatcher:
    """Keras ``fit(x, y, batch_size, shuffle=True)`` batching: a fresh permutation every
    epoch and a partial last batch (60000 = 468*128 + 96).

    With CUDA available the rows of a batch are gathered straight into a ring of page-locked buffers
    (``np.take(..., out=...)``): no per-batch allocation, and the trainer's host->device copy of the batch is
    truly asynchronous.  With ``prefetch=True`` (or ``HVD_PREFETCH_THREAD=1``) the gather (~50 us for a 128 x 784
    float batch out of 60,000 shuffled rows) runs on a thread, up to ``AHEAD`` batches in front of the consumer;
    ``np.take`` releases the GIL, so it can overlap the training thread's launch work.

    A yielded batch stays valid until ``RING - AHEAD - 1`` further batches have been taken - ``fit`` never runs
    more than 8 steps ahead of the device, so the copy of a batch has left host memory long before its slot
    is reused."""

    RING = 24
    AHEAD = 6

    def __init__(self, arrays: Sequence[np.ndarray], batch_size: int, shuffle: bool, seed=None, pinned=None,
                 prefetch=None):
        self.arrays = [np.ascontiguousarray(a) for a in arrays]
        self.n = len(self.arrays[0])
        self.batch_size = int(batch_size)
        self.shuffle = shuffle
        self.rng = np.random.default_rng(seed)
        self._ring = None
        self._k = 0
        if pinned is None:
            pinned = _cuda_available()
        if pinned:
            try:
                import torch
                self._keep = [torch.empty((self.RING, self.batch_size) + a.shape[1:],
                                          dtype=torch.from_numpy(a[:0]).dtype).pin_memory() for a in self.arrays]
                self._ring = [t.numpy() for t in self._keep]
            except Exception:                     # no CUDA runtime / unsupported dtype: plain batches
                self._ring = None
        if self._ring is None and prefetch:       # (tests) the threaded path without page-locked memory
            self._ring = [np.empty((self.RING, self.batch_size) + a.shape[1:], dtype=a.dtype) for a in self.arrays]
        if prefetch is None:
            prefetch = os.environ.get("HVD_PREFETCH_THREAD", "0") == "1"
        self.prefetch = bool(prefetch) and self._ring is not None

    @property
    def steps(self) -> int:
        return -(-self.n // self.batch_size)

    def _gather(self, idx):
        if self._ring is None:
            return tuple(a[idx] for a in self.arrays)
        k, self._k = self._k, (self._k + 1) % self.RING
        out = []
        for a, r in zip(self.arrays, self._ring):
            dst = r[k][:len(idx)]
            np.take(a, idx, axis=0, out=dst, mode="clip")      # "clip" = unbuffered; indices are valid by construction
            out.append(dst)
        return tuple(out)

    def _index_batches(self):
        order = self.rng.permutation(self.n) if self.shuffle else None
        for s in range(0, self.n, self.batch_size):
            if order is None:
                yield None if self._ring is None else np.arange(s, min(s + self.batch_size, self.n)), s
            else:
                yield order[s:s + self.batch_size], s

    def _take(self, idx, s):
        if idx is None:                           # unshuffled, no ring: plain views of the source arrays
            return tuple(a[s:s + self.batch_size] for a in self.arrays)
        return self._gather(idx)

    def epoch(self):
        if not self.prefetch:
            for idx, s in self._index_batches():
                yield self._take(idx, s)
            return
        import queue
        import threading
        q: "queue.Queue" = queue.Queue(maxsize=self.AHEAD)
        stop = threading.Event()

        def put(item) -> bool:
            while not stop.is_set():
                try:
                    q.put(item, timeout=0.05)
                    return True
                except queue.Full:
                    pass
            return False

        def produce():
            try:
                for idx, s in self._index_batches():
                    if not put(self._take(idx, s)):
                        return
                put(None)
            except BaseException as e:            # surface producer errors in the consumer
                put(e)

        t = threading.Thread(target=produce, name="hvd-batch-prefetch", daemon=True)
        t.start()
        try:
            while True:
                item = q.get()
                if item is None:
                    break
                if isinstance(item, BaseException):
                    raise item
                yield item
        finally:                                  # also runs when the consumer abandons the epoch early
            stop.set()
            t.join()


def _cuda_available() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except ImportError:  # pragma: no cover
        return False
