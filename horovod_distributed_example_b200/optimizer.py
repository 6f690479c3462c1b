"""``hvd.DistributedOptimizer`` - wraps any optimizer so gradients are averaged across
ranks before the update.

Reference use: ``opt = hvd.DistributedOptimizer(opt)`` (``mnist_keras.py:87``,
``tensorflow2_keras_mnist.py:58``).

Kept from Horovod: the wrapper is a dynamically created subclass of the
wrapped optimizer's class *with the same class name* (so a saved ``training_config``
still says ``Adadelta``), re-instantiated from ``get_config()``; it overrides
``get_gradients``; at size 1 the collective is skipped entirely.

Different from Horovod: on B200 with peer memory up, ``apply_gradients`` does not call
an allreduce and then the update - it launches ONE kernel per bucket that reads the peer
GPUs' gradient ranges over NVSwitch, sums in a fixed order, scales by 1/N and applies
the Adadelta/Adam/SGD update in registers (``csrc/comm/fused_allreduce.cu``).  Buckets
are launched on a side stream as soon as backward has produced them.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import basics, collectives
from .collectives import Compression
from .parallel import buckets as bucket_mod


class _DistributedMixin:
    _hvd_distributed = True

    def _hvd_setup(self, name, compression, sparse_as_dense, backward_passes_per_step):
        self._hvd_name = name or f"Distributed{type(self).__name__}"
        self._compression = compression
        self._sparse_as_dense = sparse_as_dense
        self._bpps = int(backward_passes_per_step)
        self._buckets: Optional[List[bucket_mod.Bucket]] = None
        self._fused = None
        self._cut_after: List[str] = []
        self._pass = 0

    # -- plan -----------------------------------------------------------------------
    def bind(self, arena) -> None:
        super().bind(arena)
        cfg = basics.config() if basics.is_initialized() else None
        thr = cfg.fusion_threshold if cfg else 64 << 20
        self._buckets = bucket_mod.plan(arena, thr, self._cut_after)
        self._plan_hash = bucket_mod.check_agreement(arena, self._buckets)
        self._fused = None
        if basics.is_initialized() and arena.device.type == "cuda" and arena.symmetric \
                and self.kind != 0:
            from .parallel import fused
            self._fused = fused.FusedReducer(arena, self, self._buckets)

    def set_bucket_cuts(self, cut_after: List[str]) -> None:
        """Force bucket boundaries after the named tensors (used for backward overlap)."""
        self._cut_after = list(cut_after)
        if self._arena is not None:
            self.bind(self._arena)

    @property
    def buckets(self):
        return self._buckets

    # -- Horovod's hook: average the gradients -----------------------------------------
    def get_gradients(self, arena) -> torch.Tensor:
        g = arena.grads
        if not basics.is_initialized() or basics.size() == 1:
            return g                       # Horovod skips the allreduce at size 1
        for b in self._buckets:
            seg = g[b.start:b.end]
            comp, ctx = self._compression.compress(seg)
            if comp is seg:
                collectives.allreduce_(seg, average=True, name=f"{self._hvd_name}_Allreduce.{b.index}")
            else:
                collectives.allreduce_(comp, average=True, name=f"{self._hvd_name}_Allreduce.{b.index}")
                seg.copy_(self._compression.decompress(comp, ctx))
        return g

    def materialize_state(self) -> None:
        """Gather sharded optimizer slots to this rank (before a checkpoint is written)."""
        if self._fused is not None:
            self._fused.materialize_state()

    def begin_step(self) -> None:
        basics.state().step += 1

    def apply_gradients(self, arena=None) -> None:
        arena = arena or self._arena
        self._pass += 1
        if self._pass % self._bpps != 0:
            return                         # local accumulation pass
        if self._fused is not None and self._compression is Compression.none:
            # fused peer-memory path: reduce + scale + update in one kernel per bucket
            self.iterations += 1
            self._fused.run_all()
            self._push_hyper()
            cfg = basics.config()
            if cfg.debug_checksum and self.iterations % cfg.debug_checksum == 0:
                self._check_replicas(arena)
            return
        super().apply_gradients(arena)


def _check_replicas_impl(arena):
    cs = int(arena.checksum())
    all_cs = collectives.allgather_object(cs)
    if any(c != all_cs[0] for c in all_cs):
        raise RuntimeError(f"replica divergence detected (HVD_DEBUG_CHECKSUM): checksums {all_cs}")


_DistributedMixin._check_replicas = staticmethod(_check_replicas_impl)


def DistributedOptimizer(optimizer, name=None, device_dense="", device_sparse="",
                         compression=Compression.none, sparse_as_dense=False,
                         backward_passes_per_step=1):
    """Return a distributed version of ``optimizer`` (same class name, same config)."""
    from .keras import optimizers as opt_mod
    optimizer = opt_mod.get(optimizer)
    if getattr(optimizer, "_hvd_distributed", False):
        raise ValueError("optimizer is already a DistributedOptimizer")
    base = type(optimizer)
    cls = type(base.__name__, (_DistributedMixin, base), {"__module__": base.__module__})
    dist_opt = cls(**optimizer.get_config())
    dist_opt.iterations = optimizer.iterations
    dist_opt._hvd_setup(name, compression, sparse_as_dense, backward_passes_per_step)
    return dist_opt
