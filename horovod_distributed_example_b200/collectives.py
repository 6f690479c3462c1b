"""Tensor-level collectives: ``allreduce``, ``broadcast``, ``allgather`` and helpers.

These are the operations Horovod's ``HorovodAllreduce/Broadcast/Allgather`` TF ops
provide underneath the reference's ``hvd.DistributedOptimizer`` (``mnist_keras.py:87``),
``BroadcastGlobalVariablesCallback`` (``mnist_keras.py:97``) and ``MetricAverageCallback``
(``tensorflow2_keras_mnist.py:77``).

Dispatch:

* size 1                        -> identity (Horovod skips the collective too);
* CUDA tensor, peer memory up   -> hand-written sm_100a peer-memory kernels
  (``parallel.peer``: one-shot / two-shot / NVLS picked per message size);
* CUDA tensor, ``HVD_ALLREDUCE_VARIANT=nccl`` -> ``torch.distributed`` NCCL (baseline);
* host tensor                   -> gloo.
"""
from __future__ import annotations

from typing import Iterable, Optional

import torch

from . import basics
from .basics import _require_init, _state, log
from .utils import nvtx

# Horovod reduce-op constants (later Horovod versions; `average=` is the 0.18 spelling).
Average = "average"
Sum = "sum"
Adasum = "adasum"  # accepted for signature parity; not implemented (not used by the reference)


class Compression:
    """Gradient compression selector (``hvd.Compression``).

    The reference calls ``DistributedOptimizer(opt)`` with the default
    ``Compression.none`` (``mnist_keras.py:87``).  ``fp16`` is provided for API parity:
    it rounds the payload to half precision on the wire and back.
    """

    class none:  # noqa: N801 - Horovod's spelling
        @staticmethod
        def compress(t):
            return t, None

        @staticmethod
        def decompress(t, ctx):
            return t

    class fp16:  # noqa: N801
        @staticmethod
        def compress(t):
            if t.dtype.is_floating_point and t.dtype != torch.float16:
                return t.to(torch.float16), t.dtype
            return t, None

        @staticmethod
        def decompress(t, ctx):
            return t.to(ctx) if ctx is not None else t


def _dist():
    import torch.distributed as dist
    return dist


def _use_peer(t: torch.Tensor) -> bool:
    if not t.is_cuda:
        return False
    cfg = _state.config
    if cfg is not None and cfg.allreduce_variant == "nccl":
        return False
    from .parallel import peer
    return _state.size > 1 and peer.available()


def _fault_skip() -> bool:
    cfg = _state.config
    f = cfg.parsed_fault() if cfg else None
    return bool(f and f[0] == _state.rank and f[1] == _state.step)


def _fault_stall() -> None:
    """HVD_FAULT_INJECT=rank:step fired on this rank: behave like a rank that hangs - stop participating (the
    peers' in-kernel waits run into their bound and name this rank), then fail.  Skipping the step and carrying on
    would not model a failure: the per-rank epoch counters would be off by one and the next step's packets would be
    taken for this one's."""
    import time
    cfg = _state.config
    wait = (cfg.barrier_abort_time() if cfg else 0.0) or 60.0
    log.error("HVD_FAULT_INJECT: rank %d stalls at step %d", _state.rank, _state.step)
    time.sleep(wait + 10.0)
    raise RuntimeError(f"fault injected: rank {_state.rank} stalled at step {_state.step}")


def allreduce_(tensor: torch.Tensor, average: bool = True, name: Optional[str] = None,
               op: Optional[str] = None) -> torch.Tensor:
    """In-place allreduce.  ``average=True`` divides by ``size()`` after the sum,
    as Horovod 0.18.1 does (sum allreduce, then a separate divide)."""
    _require_init()
    if op is not None:
        if op == Adasum:
            raise NotImplementedError("Adasum is not part of the reference's surface")
        average = (op == Average)
    if _state.size == 1:
        return tensor
    tl = _state.timeline
    if tl is not None:
        tl.begin(name or "allreduce", "ALLREDUCE")
    nvtx.push(name or "hvd.allreduce")
    try:
        if _use_peer(tensor):
            from .parallel import peer
            peer.context().allreduce_(tensor, average=average)
        else:
            dist = _dist()
            group = None if tensor.is_cuda else _state.cpu_group
            work_t = tensor if tensor.is_contiguous() else tensor.contiguous()
            dist.all_reduce(work_t, op=dist.ReduceOp.SUM, group=group)
            if average:
                if work_t.dtype.is_floating_point:
                    work_t.div_(_state.size)
                else:
                    work_t.copy_(torch.div(work_t, _state.size, rounding_mode="floor"))
            if work_t is not tensor:
                tensor.copy_(work_t)
    finally:
        nvtx.pop()
        if tl is not None:
            tl.end(name or "allreduce", "ALLREDUCE")
    return tensor


def allreduce(tensor, average: bool = True, name: Optional[str] = None,
              compression=Compression.none, op: Optional[str] = None):
    """Out-of-place allreduce of a tensor (or Python/NumPy scalar -> float)."""
    _require_init()
    if not isinstance(tensor, torch.Tensor):
        t = torch.as_tensor(tensor, dtype=torch.float64)
        out = allreduce_(t.clone(), average=average, name=name, op=op)
        return out.item() if out.ndim == 0 else out
    comp, ctx = compression.compress(tensor)
    out = allreduce_(comp.clone(), average=average, name=name, op=op)
    return compression.decompress(out, ctx)


def grouped_allreduce_(tensors: Iterable[torch.Tensor], average: bool = True) -> None:
    """Allreduce several tensors with one collective (packs, reduces, unpacks).

    Used for the metric vector (one packed allreduce instead of one
    per metric).  Parameters and gradients never go through here: they already live
    contiguously in the arena, so there is no pack/unpack copy for them.
    """
    tensors = list(tensors)
    if not tensors or _state.size == 1:
        return
    flat = torch.cat([t.reshape(-1).to(torch.float32) for t in tensors])
    allreduce_(flat, average=average)
    off = 0
    for t in tensors:
        n = t.numel()
        t.copy_(flat[off:off + n].view_as(t).to(t.dtype))
        off += n


def broadcast_(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> torch.Tensor:
    """In-place broadcast from ``root_rank``; afterwards every rank holds bit-identical data."""
    _require_init()
    if _state.size == 1:
        return tensor
    if not (0 <= root_rank < _state.size):
        raise ValueError(f"root_rank {root_rank} out of range for size {_state.size}")
    tl = _state.timeline
    if tl is not None:
        tl.begin(name or "broadcast", "BROADCAST")
    nvtx.push(name or "hvd.broadcast")
    try:
        if _use_peer(tensor):
            from .parallel import peer
            peer.context().broadcast_(tensor, root_rank)
        else:
            dist = _dist()
            group = None if tensor.is_cuda else _state.cpu_group
            if tensor.is_contiguous():
                dist.broadcast(tensor, src=root_rank, group=group)
            else:
                tmp = tensor.contiguous()
                dist.broadcast(tmp, src=root_rank, group=group)
                tensor.copy_(tmp)
    finally:
        nvtx.pop()
        if tl is not None:
            tl.end(name or "broadcast", "BROADCAST")
    return tensor


def broadcast(tensor: torch.Tensor, root_rank: int = 0, name: Optional[str] = None) -> torch.Tensor:
    return broadcast_(tensor.clone(), root_rank, name)


def allgather(tensor: torch.Tensor, name: Optional[str] = None) -> torch.Tensor:
    """Concatenate every rank's tensor along dim 0 (first dims may differ, as in Horovod)."""
    _require_init()
    if _state.size == 1:
        return tensor.clone()
    dist = _dist()
    group = None if tensor.is_cuda else _state.cpu_group
    t = tensor.contiguous()
    if t.ndim == 0:
        t = t.reshape(1)
    n0 = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    sizes = [torch.zeros_like(n0) for _ in range(_state.size)]
    dist.all_gather(sizes, n0, group=group)
    sizes = [int(s.item()) for s in sizes]
    mx = max(sizes)
    pad = t
    if t.shape[0] != mx:
        pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
    outs = [torch.empty_like(pad) for _ in range(_state.size)]
    dist.all_gather(outs, pad, group=group)
    return torch.cat([o[:s] for o, s in zip(outs, sizes)], dim=0)


def broadcast_object(obj, root_rank: int = 0):
    """Pickle-broadcast an arbitrary Python object over the host channel."""
    _require_init()
    if _state.size == 1:
        return obj
    dist = _dist()
    box = [obj if _state.rank == root_rank else None]
    dist.broadcast_object_list(box, src=root_rank, group=_state.cpu_group)
    return box[0]


def allgather_object(obj):
    _require_init()
    if _state.size == 1:
        return [obj]
    dist = _dist()
    out = [None] * _state.size
    dist.all_gather_object(out, obj, group=_state.cpu_group)
    return out


def barrier() -> None:
    _require_init()
    if _state.size > 1:
        _dist().barrier(group=_state.cpu_group)


def broadcast_variables(variables: Iterable[torch.Tensor], root_rank: int = 0) -> None:
    """Broadcast a list of tensors in place (``hvd.broadcast_variables``)."""
    for i, v in enumerate(variables):
        broadcast_(v, root_rank, name=f"broadcast.{i}")


def broadcast_parameters(params, root_rank: int = 0) -> None:
    """``hvd.broadcast_parameters`` for a ``state_dict`` or named-parameter iterable."""
    items = params.items() if isinstance(params, dict) else params
    for name, p in items:
        if isinstance(p, torch.Tensor):
            broadcast_(p.data, root_rank, name=f"broadcast.{name}")
