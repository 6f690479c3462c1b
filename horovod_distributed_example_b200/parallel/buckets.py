"""Static gradient-bucket plan.

Horovod negotiates which tensors to fuse every 5 ms cycle (coordinator protocol,
response cache, fusion-buffer manager - SURVEY.md section 2.4).  Every rank here runs the
same model, so the plan is computed ONCE from the arena layout, its hash is checked
across ranks at optimizer construction, and the data path carries no control traffic.

A bucket is a contiguous element range ``[start, end)`` of the gradient slab that covers
whole tensors (the arena is already in backward-ready order).  Buckets close when adding
the next tensor would exceed ``HVD_FUSION_THRESHOLD`` bytes, or at explicit cut points
(the MNIST engine cuts after ``dense_1/kernel`` so 98% of the bytes start reducing while
the conv backward is still running).

The planner itself is native (``csrc/core/core.cpp: hvd_plan_buckets`` via ``libhvd_core.so``); the
pure-Python twin below is the oracle for its unit test and the fallback when the
library has not been built.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence

from .arena import ALIGN, ParamArena


@dataclass(frozen=True)
class Bucket:
    index: int
    start: int          # element offset in the slab
    end: int            # exclusive, ALIGN-rounded
    first_tensor: int
    last_tensor: int    # inclusive

    @property
    def numel(self) -> int:
        return self.end - self.start

    @property
    def nbytes(self) -> int:
        return self.numel * 4


def _fnv1a(h: int, data: bytes) -> int:
    for b in data:
        h ^= b
        h = (h * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return h


def plan_python(offsets: Sequence[int], numels: Sequence[int], slab_elems: int,
                threshold_bytes: int, cuts: Sequence[int] = ()) -> List[Bucket]:
    """Reference implementation (mirrors ``hvd_plan_buckets``)."""
    n = len(offsets)
    cutset = set(int(c) for c in cuts)
    out: List[Bucket] = []
    first = 0
    while first < n:
        last = first
        start = offsets[first]
        while last + 1 < n:
            if last in cutset:
                break
            nxt_end = offsets[last + 1] + numels[last + 1]
            if (nxt_end - start) * 4 > threshold_bytes:
                break
            last += 1
        end = offsets[last + 1] if last + 1 < n else slab_elems
        out.append(Bucket(len(out), start, end, first, last))
        first = last + 1
    return out


def plan_hash(names: Sequence[str], shapes: Sequence[Sequence[int]], buckets: Sequence[Bucket]) -> int:
    # Shapes and order only: layer names are auto-numbered per process (a rank that built another
    # model first - e.g. rank 0 reloading a checkpoint - would otherwise disagree for no reason).
    h = 0xCBF29CE484222325
    for _nme, shp in zip(names, shapes):
        h = _fnv1a(h, (",".join(str(int(s)) for s in shp) + ";").encode())
    for b in buckets:
        h = _fnv1a(h, f"{b.start}:{b.end}|".encode())
    return h


def plan(arena: ParamArena, threshold_bytes: int, cut_after: Sequence[str] = ()) -> List[Bucket]:
    offsets = [s.offset for s in arena.specs]
    numels = [s.numel for s in arena.specs]
    names = [s.name for s in arena.specs]
    cuts = [names.index(c) for c in cut_after if c in names]
    from ..ops import native
    lib = native.core_lib(required=False)
    if lib is None:
        return plan_python(offsets, numels, arena.n, threshold_bytes, cuts)
    n = len(offsets)
    I64 = ctypes.c_int64
    off_a = (I64 * n)(*offsets)
    num_a = (I64 * n)(*numels)
    cut_a = (I64 * max(len(cuts), 1))(*cuts)
    out = (I64 * (4 * n))()
    nb = lib.hvd_plan_buckets(off_a, num_a, n, I64(arena.n), I64(threshold_bytes),
                              cut_a, len(cuts), out)
    if nb < 0:
        raise RuntimeError("hvd_plan_buckets failed")
    return [Bucket(i, out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]) for i in range(nb)]


def check_agreement(arena: ParamArena, buckets: Sequence[Bucket]) -> int:
    """Allgather the plan hash over the host channel; raise if any rank disagrees.

    This is the one-time replacement for Horovod's per-step negotiation."""
    from .. import basics, collectives
    h = plan_hash([s.name for s in arena.specs], [s.shape for s in arena.specs], buckets)
    if basics.is_initialized() and basics.size() > 1:
        all_h = collectives.allgather_object(h)
        if any(x != h for x in all_h):
            bad = [i for i, x in enumerate(all_h) if x != all_h[0]]
            raise RuntimeError(
                f"bucket plan mismatch across ranks (ranks {bad} differ from rank 0): every rank "
                "must build the same model in the same order")
    return h
