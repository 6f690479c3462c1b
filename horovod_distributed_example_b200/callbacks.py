"""``hvd.callbacks``: BroadcastGlobalVariables, MetricAverage, LearningRateSchedule,
LearningRateWarmup.

Reference use: ``BroadcastGlobalVariablesCallback(0)`` (``mnist_keras.py:97``,
``tensorflow2_keras_mnist.py:71``), ``MetricAverageCallback()``
(``tensorflow2_keras_mnist.py:77``), ``LearningRateWarmupCallback(warmup_epochs=3,
verbose=1)`` (``tensorflow2_keras_mnist.py:82``).  Behaviour per SURVEY.md F7-F9.
"""
from __future__ import annotations

import torch

from . import basics, collectives
from .keras.callbacks import Callback


class BroadcastGlobalVariablesCallback(Callback):
    """Copy rank ``root_rank``'s model weights *and optimizer state* to every rank once,
    so all workers start identical - from random init or from a checkpoint restored on
    rank 0 (``mnist_keras.py:94-96``).

    Horovod 0.18.1 issues one broadcast per variable (25+ collectives for this model).
    Params and optimizer slots are contiguous in the arena here, so it is one broadcast
    of the params slab and one per slot slab - on B200 a single peer-memory pull kernel
    per slab - plus one tiny host broadcast for the iteration counter and LR.

    Fires at ``on_train_begin`` (the arena and its slots exist by then; there is no lazy
    slot creation to wait for, which is why later Horovod versions moved it to the end
    of the first batch)."""

    def __init__(self, root_rank: int = 0, device: str = ""):
        super().__init__()
        self.root_rank = root_rank
        self.broadcast_done = False

    def on_train_begin(self, logs=None):
        if self.broadcast_done:
            return
        if basics.size() > 1:
            m = self.model
            m._sync_backend()
            arena = m.arena
            opt = m.optimizer
            if basics.rank() == self.root_rank and opt is not None and hasattr(opt, "materialize_state"):
                # optimizer state of owner-updated ranges lives with its owners: the root completes its copy
                # (one-sided peer reads) before it becomes everybody's state
                opt.materialize_state()
            with torch.no_grad():
                collectives.broadcast_(arena.params, self.root_rank, name="broadcast.params")
                for i, s in enumerate(arena.slots):
                    collectives.broadcast_(s, self.root_rank, name=f"broadcast.slot{i}")
            opt = m.optimizer
            if opt is not None:
                it, lr = collectives.broadcast_object((opt.iterations, opt.lr), self.root_rank)
                opt.iterations = it
                opt.lr = lr
            be = getattr(m, "_backend", None)
            if be is not None and hasattr(be, "weights_changed"):
                be.weights_changed()
        self.broadcast_done = True


class MetricAverageCallback(Callback):
    """At epoch end replace every metric in ``logs`` (in place) with its average over
    ranks, so later callbacks see the averaged values.  Metrics are visited in sorted
    order on every rank; they are packed into ONE vector and reduced with one collective
    (Horovod does one scalar allreduce per metric - SURVEY.md K-C3)."""

    def __init__(self, device: str = ""):
        super().__init__()

    def on_epoch_end(self, epoch, logs=None):
        if logs is None or basics.size() == 1:
            return
        keys = sorted(k for k, v in logs.items() if _is_number(v))
        if not keys:
            return
        vec = torch.tensor([float(logs[k]) for k in keys], dtype=torch.float64)
        dev = basics.device()
        if dev.type == "cuda":
            # the metric vector rides the same peer-memory allreduce kernels as the gradients
            vec = vec.to(dev)
        collectives.allreduce_(vec, average=True, name="metric_average")
        for k, v in zip(keys, vec.tolist()):
            logs[k] = v


def _is_number(v) -> bool:
    if isinstance(v, (int, float)):
        return True
    if isinstance(v, torch.Tensor):
        return v.numel() == 1
    try:
        float(v)
        return True
    except (TypeError, ValueError):
        return False


class LearningRateScheduleCallback(Callback):
    """Multiply the initial LR by ``multiplier(epoch)`` between ``start_epoch`` and
    ``end_epoch``.  With ``staircase=False`` the epoch passed to ``multiplier`` is
    fractional and the LR is adjusted every batch.  ``momentum_correction`` scales a
    ``momentum`` attribute by ``new_lr/old_lr`` for the batch and restores it after
    (arXiv:1706.02677)."""

    def __init__(self, multiplier, start_epoch=0, end_epoch=None, staircase=True,
                 momentum_correction=True, steps_per_epoch=None, initial_lr=None):
        super().__init__()
        self.start_epoch = start_epoch
        self.end_epoch = end_epoch
        self.staircase = staircase
        self.momentum_correction = momentum_correction
        self.initial_lr = initial_lr
        self.restore_momentum = None
        self.steps_per_epoch = steps_per_epoch
        self.current_epoch = None
        if not callable(multiplier):
            self.staircase = True
            self.multiplier = lambda epoch: multiplier
        else:
            self.multiplier = multiplier

    def _autodetect_steps_per_epoch(self):
        if self.params.get("steps"):
            return self.params["steps"]
        if self.params.get("samples") and self.params.get("batch_size"):
            return self.params["samples"] // self.params["batch_size"]
        raise ValueError("Could not autodetect the number of steps per epoch. "
                         "Please specify the steps_per_epoch parameter.")

    def _adjust_learning_rate(self, epoch):
        opt = self.model.optimizer
        old_lr = opt.lr
        new_lr = self.initial_lr * self.multiplier(epoch)
        opt.lr = new_lr
        if hasattr(opt, "momentum") and self.momentum_correction and old_lr > 0:
            self.restore_momentum = opt.momentum
            opt.momentum = opt.momentum * new_lr / old_lr
            opt._hyper_changed()

    def _restore_momentum_if_needed(self):
        if self.restore_momentum is not None:
            self.model.optimizer.momentum = self.restore_momentum
            self.model.optimizer._hyper_changed()
            self.restore_momentum = None

    def on_train_begin(self, logs=None):
        if self.initial_lr is None:
            self.initial_lr = self.model.optimizer.lr
        if not self.staircase and not self.steps_per_epoch:
            self.steps_per_epoch = self._autodetect_steps_per_epoch()

    def on_epoch_begin(self, epoch, logs=None):
        self.current_epoch = epoch

    def on_batch_begin(self, batch, logs=None):
        if (self.current_epoch < self.start_epoch or
                (self.end_epoch is not None and self.current_epoch >= self.end_epoch)):
            return                                     # outside the adjustment scope
        if self.staircase and batch == 0:
            self._adjust_learning_rate(self.current_epoch)
        elif not self.staircase:
            epoch = self.current_epoch + float(batch) / self.steps_per_epoch
            self._adjust_learning_rate(epoch)

    def on_batch_end(self, batch, logs=None):
        self._restore_momentum_if_needed()

    def on_epoch_end(self, epoch, logs=None):
        if logs is not None:
            logs["lr"] = self.model.optimizer.lr    # written after MetricAverage ran -> not averaged


class LearningRateWarmupCallback(LearningRateScheduleCallback):
    """Ramp the LR linearly from ``initial_lr/size`` to ``initial_lr`` over
    ``warmup_epochs``, adjusting every batch::

        e  = epoch + batch/steps_per_epoch + 1/steps_per_epoch
        lr = initial_lr * (1/size) * (e*(size-1)/warmup_epochs + 1)

    The script has already scaled the LR by ``size`` (``tensorflow2_keras_mnist.py:55``), so
    this reaches exactly that value at the end of epoch ``warmup_epochs``.  At size 1 the
    multiplier is 1 throughout."""

    def __init__(self, warmup_epochs=5, momentum_correction=True, steps_per_epoch=None,
                 verbose=0, initial_lr=None):
        def multiplier(epoch):
            # ``epoch`` is fractional; +1/steps so the last batch of the warmup hits 1.0
            epoch += 1.0 / self.steps_per_epoch
            n = basics.size()
            return 1.0 / n * (epoch * (n - 1) / warmup_epochs + 1)

        super().__init__(multiplier, start_epoch=0, end_epoch=warmup_epochs, staircase=False,
                         momentum_correction=momentum_correction,
                         steps_per_epoch=steps_per_epoch, initial_lr=initial_lr)
        self.warmup_epochs = warmup_epochs
        self.verbose = verbose

    def on_epoch_end(self, epoch, logs=None):
        super().on_epoch_end(epoch, logs)
        if epoch == self.end_epoch - 1 and self.verbose > 0:
            print("\nEpoch %d: finished gradual learning rate warmup to %g." %
                  (epoch + 1, self.model.optimizer.lr))
