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
        self._hooked = False

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
        pipe = getattr(self, "_pipe", None)
        if pipe is not None:
            pipe.materialize_state()           # pipelined engine: the big kernel's state rows live with their owners
        elif self._fused is not None:
            self._fused.materialize_state()

    def begin_step(self) -> None:
        basics.state().step += 1
        if self._hooked:
            self._pending = list(self._bucket_size)
            self._launched = [False] * len(self._buckets)

    # -- backward overlap for the torch-autograd backend ---------------------------------------
    def attach_model(self, model) -> None:
        """Overlap communication with backward (what Horovod gets from its background thread and DDP
        from its reducer): a post-accumulate-grad hook per parameter counts down its bucket; the
        moment a bucket's last gradient lands, its fused allreduce+update kernel is launched on a
        side stream behind an event, while autograd keeps producing the earlier layers' gradients."""
        self._hooked = False
        if self._fused is None or len(self._buckets) < 2 or getattr(model, "_named", None) is None:
            return
        arena = self._arena
        index = {s.name: i for i, s in enumerate(arena.specs)}
        owner = {}
        for b in self._buckets:
            for t in range(b.first_tensor, b.last_tensor + 1):
                owner[t] = b.index
        self._bucket_size = [b.last_tensor - b.first_tensor + 1 for b in self._buckets]
        self._pending = list(self._bucket_size)
        self._launched = [False] * len(self._buckets)
        for h in getattr(self, "_hook_handles", []):
            h.remove()
        self._hook_handles = []
        for name, p in model._named:
            bi = owner[index[name]]
            self._hook_handles.append(p.register_post_accumulate_grad_hook(self._make_hook(bi)))
        self._hooked = True

    def _make_hook(self, bi: int):
        def hook(_param):
            self._pending[bi] -= 1
            if self._pending[bi] == 0 and self._pass % self._bpps == self._bpps - 1:
                self._launch_bucket(bi)
        return hook

    def _launch_bucket(self, bi: int) -> None:
        if self._launched[bi]:
            return
        self._launched[bi] = True
        f = self._fused
        ev = f._events[bi]
        ev.record(torch.cuda.current_stream(self._arena.device))
        f.side_stream.wait_event(ev)
        f.run_bucket(bi, 0, 1, stream=f.side_stream)

    def apply_gradients(self, arena=None) -> None:
        arena = arena or self._arena
        self._pass += 1
        if self._pass % self._bpps != 0:
            return                         # local accumulation pass
        if self._fused is not None and self._compression is Compression.none:
            # fused peer-memory path: reduce + scale + update in one kernel per bucket
            self._fused.ctx.check_error()      # host watchdog: a barrier that timed out names the stalled peer
            if collectives._fault_skip():      # HVD_FAULT_INJECT=rank:step - this rank hangs, then fails
                collectives._fault_stall()
            self.iterations += 1
            if self._hooked:
                for bi in range(len(self._buckets)):      # anything the hooks did not launch (unused params)
                    self._launch_bucket(bi)
                torch.cuda.current_stream(arena.device).wait_stream(self._fused.side_stream)
            else:
                self._fused.run_all()
            self._push_hyper()
            self._debug_after_step(arena, 0)
            return
        super().apply_gradients(arena)
        self._debug_after_step(arena, 0)

    def _debug_after_step(self, arena, which: int = 0) -> None:
        """replica checksum every HVD_DEBUG_CHECKSUM steps; HVD_DEBUG_POISON=1 NaN-fills
        the gradient ranges the update has just consumed, so a later kernel that reads a gradient nobody
        rewrote (stale data that would otherwise look plausible) turns the loss into NaN at once."""
        cfg = basics.config()
        if cfg.debug_checksum and self.iterations % cfg.debug_checksum == 0:
            self._check_replicas(arena)
        if cfg.debug_poison:
            arena.poison_grads(which)


def _check_replicas_impl(arena):
    torch.cuda.synchronize(arena.device) if arena.device.type == "cuda" else None
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
