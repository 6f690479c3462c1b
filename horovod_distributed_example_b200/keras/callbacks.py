"""Keras-style training callbacks: base class, History, progress bar, ModelCheckpoint,
TensorBoard.

Reference use: rank-0-only ``ModelCheckpoint('checkpoint-{epoch}.h5')`` and
``TensorBoard(log_dir, update_freq='batch')`` (``mnist_keras.py:101-105``,
``tensorflow2_keras_mnist.py:86-89``), ``verbose=1`` only on rank 0
(``mnist_keras.py:111``, ``tensorflow2_keras_mnist.py:92``).

Log values may be 0-dim device tensors; a callback calls ``float(v)`` only when it
really needs the number, so nothing here forces a device sync every step unless the
user asked for per-batch output (``update_freq='batch'`` does, as in the reference).
"""
from __future__ import annotations

import collections
import os
import sys
import time
from typing import Dict, List, Optional


class Callback:
    def __init__(self):
        self.model = None
        self.params: Dict = {}

    def set_model(self, model):
        self.model = model

    def set_params(self, params):
        self.params = params

    def on_train_begin(self, logs=None): ...
    def on_train_end(self, logs=None): ...
    def on_epoch_begin(self, epoch, logs=None): ...
    def on_epoch_end(self, epoch, logs=None): ...
    def on_batch_begin(self, batch, logs=None): ...
    def on_batch_end(self, batch, logs=None): ...
    # tf.keras spellings
    def on_train_batch_begin(self, batch, logs=None):
        self.on_batch_begin(batch, logs)

    def on_train_batch_end(self, batch, logs=None):
        self.on_batch_end(batch, logs)


class CallbackList:
    """Calls callbacks *in list order* - the order matters: MetricAverage must run before
    any metrics-consuming callback (``tensorflow2_keras_mnist.py:73-76``)."""

    def __init__(self, callbacks: List[Callback], model, params):
        self.callbacks = list(callbacks)
        for c in self.callbacks:
            c.set_model(model)
            c.set_params(params)

    def _call(self, hook, *a):
        for c in self.callbacks:
            getattr(c, hook)(*a)

    def on_train_begin(self, logs=None): self._call("on_train_begin", logs)
    def on_train_end(self, logs=None): self._call("on_train_end", logs)
    def on_epoch_begin(self, e, logs=None): self._call("on_epoch_begin", e, logs)
    def on_epoch_end(self, e, logs=None): self._call("on_epoch_end", e, logs)
    def on_batch_begin(self, b, logs=None): self._call("on_train_batch_begin", b, logs)
    def on_batch_end(self, b, logs=None): self._call("on_train_batch_end", b, logs)


class History(Callback):
    def on_train_begin(self, logs=None):
        self.epoch: List[int] = []
        self.history: Dict[str, List[float]] = {}

    def on_epoch_end(self, epoch, logs=None):
        self.epoch.append(epoch)
        for k, v in (logs or {}).items():
            self.history.setdefault(k, []).append(float(v))


class ProgbarLogger(Callback):
    """``verbose=1`` progress output.  Prints at most every ``interval`` seconds so the
    device is not synced every step."""

    def __init__(self, interval: float = 0.5, stream=None):
        super().__init__()
        self.interval = interval
        self.stream = stream or sys.stdout

    def on_epoch_begin(self, epoch, logs=None):
        self.stream.write(f"Epoch {epoch + 1}/{self.params.get('epochs', '?')}\n")
        self._t0 = self._last = time.time()
        self._seen = 0

    def _fmt(self, logs):
        return " - ".join(f"{k}: {float(v):.4f}" for k, v in logs.items()
                          if k not in ("batch", "size"))

    def on_batch_end(self, batch, logs=None):
        self._seen = batch + 1
        now = time.time()
        if now - self._last >= self.interval:
            self._last = now
            steps = self.params.get("steps")
            self.stream.write(f"\r{self._seen}/{steps if steps else '?'} - {self._fmt(logs or {})}")
            self.stream.flush()

    def on_epoch_end(self, epoch, logs=None):
        steps = self.params.get("steps") or self._seen
        dt = time.time() - self._t0
        self.stream.write(f"\r{steps}/{steps} - {dt:.0f}s - {self._fmt(logs or {})}\n")
        self.stream.flush()


class ModelCheckpoint(Callback):
    """Keras defaults: ``save_weights_only=False, save_best_only=False, period=1`` - a
    full-model HDF5 file every epoch, 1-based epoch number in the name
    (``checkpoint-1.h5`` first)."""

    def __init__(self, filepath, monitor="val_loss", verbose=0, save_best_only=False,
                 save_weights_only=False, mode="auto", period=1):
        super().__init__()
        self.filepath = filepath
        self.monitor = monitor
        self.verbose = verbose
        self.save_best_only = save_best_only
        self.save_weights_only = save_weights_only
        self.period = period
        self._since = 0
        if mode == "max" or (mode == "auto" and ("acc" in monitor or monitor.startswith("fmeasure"))):
            self._better, self.best = (lambda a, b: a > b), -float("inf")
        else:
            self._better, self.best = (lambda a, b: a < b), float("inf")

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        self._since += 1
        if self._since < self.period:
            return
        self._since = 0
        path = self.filepath.format(epoch=epoch + 1, **{k: float(v) for k, v in logs.items()})
        if self.save_best_only:
            cur = logs.get(self.monitor)
            if cur is None or not self._better(float(cur), self.best):
                return
            self.best = float(cur)
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        if self.save_weights_only:
            self.model.save_weights(path)
        else:
            self.model.save(path)
        if self.verbose:
            print(f"\nEpoch {epoch + 1:05d}: saving model to {path}")


class TensorBoard(Callback):
    """Writes TF event files (own writer, ``utils/tfevents.py``) with the same scalar tags
    Keras uses: per-batch ``batch_loss``/``batch_<metric>`` when ``update_freq='batch'``,
    per-epoch ``epoch_loss``, ``epoch_<metric>``, ``epoch_val_*`` and ``epoch_lr``.

    Every scalar is mirrored to the metrics sink (``utils/metrics_sink.py``), which is the
    local stand-in for ``gradient_utils.metrics.init(sync_tensorboard=True)``
    (``mnist_keras.py:22-23``)."""

    def __init__(self, log_dir="./logs", update_freq="epoch", histogram_freq=0, write_graph=False,
                 **_ignored):
        super().__init__()
        self.log_dir = log_dir
        self.update_freq = update_freq
        self._writer = None
        self._val_writer = None
        self._step = 0
        self._pending = collections.deque()      # (writer, tag, value, step) whose value is still on its way

    def _w(self):
        if self._writer is None:
            from ..utils.tfevents import EventWriter
            self._writer = EventWriter(os.path.join(self.log_dir, "train"))
        return self._w_cached()

    def _w_cached(self):
        return self._writer

    def _vw(self):
        if self._val_writer is None:
            from ..utils.tfevents import EventWriter
            self._val_writer = EventWriter(os.path.join(self.log_dir, "validation"))
        return self._val_writer

    def _emit(self, writer, tag, value, step):
        from ..utils import metrics_sink
        v = float(value)
        writer.add_scalar(tag, v, step)
        metrics_sink.record(tag, v, step)

    def _drain(self, force: bool):
        """Write the per-batch scalars whose values have arrived (all of them when ``force``), in order.
        With an asynchronous backend a batch's values land while the next batch is already running, so they
        are written one batch late instead of stalling the training loop on every step."""
        from .utils import is_ready
        while self._pending and (force or is_ready(self._pending[0][2])):
            self._emit(*self._pending.popleft())

    def on_batch_end(self, batch, logs=None):
        self._drain(False)
        self._step += 1
        if self.update_freq == "batch" or (isinstance(self.update_freq, int)
                                           and self._step % self.update_freq == 0):
            w = self._w()
            for k, v in (logs or {}).items():
                if k in ("batch", "size"):
                    continue
                self._pending.append((w, "batch_" + k, v, self._step))
            self._drain(False)

    def on_epoch_end(self, epoch, logs=None):
        self._drain(True)
        for k, v in (logs or {}).items():
            if k.startswith("val_"):
                self._emit(self._vw(), "epoch_" + k[4:], v, epoch)
            else:
                self._emit(self._w(), "epoch_" + k, v, epoch)
        self.flush()

    def flush(self):
        for w in (self._writer, self._val_writer):
            if w is not None:
                w.flush()

    def on_train_end(self, logs=None):
        self._drain(True)
        for w in (self._writer, self._val_writer):
            if w is not None:
                w.close()
        self._writer = self._val_writer = None


class LambdaCallback(Callback):
    def __init__(self, **hooks):
        super().__init__()
        for k, fn in hooks.items():
            if fn is not None:
                setattr(self, k, fn)


class EarlyStopping(Callback):
    def __init__(self, monitor="val_loss", min_delta=0.0, patience=0, mode="min"):
        super().__init__()
        self.monitor, self.min_delta, self.patience = monitor, min_delta, patience
        self.sign = 1.0 if mode == "min" else -1.0

    def on_train_begin(self, logs=None):
        self.best, self.wait = float("inf"), 0

    def on_epoch_end(self, epoch, logs=None):
        cur = (logs or {}).get(self.monitor)
        if cur is None:
            return
        cur = self.sign * float(cur)
        if cur < self.best - self.min_delta:
            self.best, self.wait = cur, 0
        else:
            self.wait += 1
            if self.wait > self.patience:
                self.model.stop_training = True


class ReduceLROnPlateau(Callback):
    """Included because the reference's comment names it as a metrics-consuming callback
    that must come after MetricAverage (``tensorflow2_keras_mnist.py:75``)."""

    def __init__(self, monitor="val_loss", factor=0.1, patience=10, min_lr=0.0, min_delta=1e-4):
        super().__init__()
        self.monitor, self.factor, self.patience = monitor, factor, patience
        self.min_lr, self.min_delta = min_lr, min_delta

    def on_train_begin(self, logs=None):
        self.best, self.wait = float("inf"), 0

    def on_epoch_end(self, epoch, logs=None):
        cur = (logs or {}).get(self.monitor)
        if cur is None:
            return
        cur = float(cur)
        if cur < self.best - self.min_delta:
            self.best, self.wait = cur, 0
            return
        self.wait += 1
        if self.wait >= self.patience:
            opt = self.model.optimizer
            opt.lr = max(opt.lr * self.factor, self.min_lr)
            self.wait = 0


class LearningRateScheduler(Callback):
    """``schedule(epoch)`` or ``schedule(epoch, lr)`` -> new learning rate, applied at the start of every epoch
    (the fused update kernels read the LR from a device block, so this costs one small upload, no re-capture)."""

    def __init__(self, schedule, verbose=0):
        super().__init__()
        self.schedule, self.verbose = schedule, verbose

    def on_epoch_begin(self, epoch, logs=None):
        opt = self.model.optimizer
        try:
            lr = self.schedule(epoch, float(opt.lr))
        except TypeError:
            lr = self.schedule(epoch)
        opt.lr = float(lr)
        if self.verbose:
            print(f"\nEpoch {epoch + 1:05d}: LearningRateScheduler setting learning rate to {opt.lr}.")

    def on_epoch_end(self, epoch, logs=None):
        if logs is not None:
            logs["lr"] = float(self.model.optimizer.lr)


class CSVLogger(Callback):
    """One row of epoch metrics per epoch (``epoch,<sorted metric names>``)."""

    def __init__(self, filename, separator=",", append=False):
        super().__init__()
        self.filename, self.sep, self.append = filename, separator, append
        self._f = None
        self._keys = None

    def on_train_begin(self, logs=None):
        new = not (self.append and os.path.exists(self.filename) and os.path.getsize(self.filename) > 0)
        self._f = open(self.filename, "a" if self.append else "w")
        self._header_due = new

    def on_epoch_end(self, epoch, logs=None):
        logs = logs or {}
        if self._keys is None:
            self._keys = sorted(logs)
        if self._header_due:
            self._f.write(self.sep.join(["epoch"] + self._keys) + "\n")
            self._header_due = False
        self._f.write(self.sep.join([str(epoch)] + [repr(float(logs[k])) if k in logs else "NA" for k in self._keys]) + "\n")
        self._f.flush()

    def on_train_end(self, logs=None):
        if self._f is not None:
            self._f.close()
            self._f = None


class TerminateOnNaN(Callback):
    """Stop training when the loss of a batch is NaN or infinite.  Checks the batches whose results have already
    arrived (no device sync is forced per step) and everything that is left at the end of the epoch."""

    def __init__(self):
        super().__init__()
        self._pending = collections.deque()

    def _check(self, force):
        import math
        from .utils import is_ready
        while self._pending and (force or is_ready(self._pending[0][1])):
            batch, v = self._pending.popleft()
            v = float(v)
            if math.isnan(v) or math.isinf(v):
                print(f"Batch {batch}: Invalid loss, terminating training")
                self.model.stop_training = True
                self._pending.clear()

    def on_batch_end(self, batch, logs=None):
        if logs and "loss" in logs:
            self._pending.append((batch, logs["loss"]))
        self._check(False)

    def on_epoch_end(self, epoch, logs=None):
        self._check(True)
