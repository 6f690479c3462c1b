"""The parameter arena: one flat fp32 buffer per rank holding params, grads and optimizer slots.

This replaces Horovod's copy-in/copy-out fusion buffer (SURVEY.md section 2.4,
``fusion_buffer_manager.cc``): parameters, gradients and optimizer state *live* in the
arena as views, so a gradient bucket is a contiguous range with no pack/unpack copies,
and - when the arena is allocated from peer-mapped symmetric memory - the collective
kernels read and write peer GPUs' ranges directly over NVSwitch.

Layout (all fp32, every tensor start aligned to ``ALIGN`` elements)::

    [ params (n) | grads A (n) | grads B (n, optional) | slot 0 (n) | slot 1 (n) | ... ]

Tensors are laid out in *backward-ready order* (last layer first): the gradients that
finish first occupy the front of the arena, so a prefix of the arena can be reduced
while backward is still producing the rest (reference call stack: SURVEY.md section 3.2).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

ALIGN = 64  # elements (256 B): float4 + TMA base alignment, and keeps shards 128B-aligned


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


@dataclass
class TensorSpec:
    name: str
    shape: Tuple[int, ...]
    offset: int = 0          # element offset inside a slab

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= int(s)
        return n


class ParamArena:
    """Flat storage for a model's trainable tensors.

    ``specs`` must be given in backward-ready order.  ``allocator(nbytes)`` may return a
    CUDA tensor carved from symmetric memory; by default ``torch.zeros`` is used.
    """

    def __init__(self, specs: Sequence[Tuple[str, Tuple[int, ...]]], device,
                 n_slots: int = 0, double_grads: bool = False, allocator=None):
        self.device = torch.device(device)
        self.specs: List[TensorSpec] = []
        off = 0
        for name, shape in specs:
            ts = TensorSpec(name, tuple(int(s) for s in shape), off)
            self.specs.append(ts)
            off = _round_up(off + ts.numel, ALIGN)
        self.n = max(off, ALIGN)                 # elements per slab (padded)
        self.n_slots = n_slots
        self.n_grad_bufs = 2 if double_grads else 1
        self.total = self.n * (1 + self.n_grad_bufs + n_slots)
        if allocator is not None:
            self.buffer = allocator(self.total * 4)
            if self.buffer.numel() < self.total:
                raise RuntimeError("arena allocator returned a short buffer")
            self.buffer = self.buffer[: self.total]
            self.buffer.zero_()
            self.symmetric = True
        else:
            self.buffer = torch.zeros(self.total, dtype=torch.float32, device=self.device)
            self.symmetric = False
        n = self.n
        # Slabs alias the buffer's storage but are NOT autograd views of it: each gets its own
        # version counter, so accumulating into the grad slab during backward does not
        # invalidate the parameter views autograd saved during forward.
        self.params = self._slab(0)
        self.grad_bufs = [self._slab((1 + i) * n) for i in range(self.n_grad_bufs)]
        base = (1 + self.n_grad_bufs) * n
        self.slots = [self._slab(base + i * n) for i in range(n_slots)]
        self._by_name: Dict[str, TensorSpec] = {s.name: s for s in self.specs}
        if len(self._by_name) != len(self.specs):
            raise ValueError("duplicate tensor names in arena")

    def _slab(self, elem_offset: int) -> torch.Tensor:
        t = torch.empty(0, dtype=torch.float32, device=self.buffer.device)
        t.set_(self.buffer.untyped_storage(), self.buffer.storage_offset() + elem_offset,
               (self.n,), (1,))
        return t

    # element offsets of each slab inside ``buffer`` (used by the native kernels)
    @property
    def param_offset(self) -> int:
        return 0

    def grad_offset(self, which: int = 0) -> int:
        return (1 + which) * self.n

    def slot_offset(self, i: int) -> int:
        return (1 + self.n_grad_bufs + i) * self.n

    @property
    def grads(self) -> torch.Tensor:
        return self.grad_bufs[0]

    def spec(self, name: str) -> TensorSpec:
        return self._by_name[name]

    def _view(self, slab: torch.Tensor, name: str) -> torch.Tensor:
        s = self._by_name[name]
        return slab[s.offset: s.offset + s.numel].view(s.shape)

    def param(self, name: str) -> torch.Tensor:
        return self._view(self.params, name)

    def grad(self, name: str, which: int = 0) -> torch.Tensor:
        return self._view(self.grad_bufs[which], name)

    def slot(self, i: int, name: str) -> torch.Tensor:
        return self._view(self.slots[i], name)

    def n_params(self) -> int:
        return sum(s.numel for s in self.specs)

    def checksum(self) -> torch.Tensor:
        """Order-sensitive integer checksum of the parameter slab (debug replica check).

        Bitwise: reinterprets the fp32 words as int32 and sums with wraparound, so two
        ranks agree iff their parameter bits agree (up to a 2^-32 collision chance)."""
        w = self.params.view(torch.int32).to(torch.int64)
        idx = torch.arange(1, w.numel() + 1, device=w.device, dtype=torch.int64)
        return ((w * (idx % 8191 + 1)).sum()) & 0x7FFFFFFFFFFF

    def poison_grads(self, which: int = 0) -> None:
        """NaN-fill every tensor's range of gradient buffer ``which`` (debug; padding is left alone because
        the update kernels run over it and the parameter checksum covers it)."""
        buf = self.grad_bufs[which]
        for s in self.specs:
            buf[s.offset:s.offset + s.numel].fill_(float("nan"))

    def resize_slots(self, n_slots: int, allocator=None, double_grads: Optional[bool] = None) -> "ParamArena":
        """Return a new arena with ``n_slots`` optimizer slots (optionally in memory from ``allocator``),
        carrying params and existing slots over."""
        if double_grads is None:
            double_grads = self.n_grad_bufs == 2
        new = ParamArena([(s.name, s.shape) for s in self.specs], self.device, n_slots,
                         double_grads=double_grads, allocator=allocator)
        new.params.copy_(self.params)
        for i in range(min(n_slots, self.n_slots)):
            new.slots[i].copy_(self.slots[i])
        return new
