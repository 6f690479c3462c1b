"""Per-bucket fused allreduce + 1/N scale + optimizer update over the symmetric arena.

This is what ``hvd.DistributedOptimizer`` runs on B200 instead of "allreduce, then update"
(reference ``mnist_keras.py:87``; Horovod path in SURVEY.md section 3.2).  One kernel launch
per bucket (``csrc/comm/fused_allreduce.cu``); the variant (one-shot / two-shot / NVLS) is
fixed per bucket at construction from the variant table, because two-shot/NVLS buckets keep
their optimizer state *sharded* by owner rank and must not change owner mid-training.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ..utils import nvtx

from .. import basics
from . import peer as peer_mod
from .peer import HvdFusedArgs, NVLS, ONESHOT, TWOSHOT, VARIANT_NAMES


class FusedReducer:
    def __init__(self, arena, optimizer, buckets, ctx: Optional[peer_mod.PeerContext] = None,
                 variants: Optional[List[int]] = None):
        self.arena, self.opt, self.buckets = arena, optimizer, list(buckets)
        self.ctx = ctx or peer_mod.context()
        buf, off = self.ctx.buffer_of(arena.buffer)
        if buf is None:
            raise RuntimeError("the arena does not live in symmetric memory")
        self.buf, self.base_off = buf, off
        if len(self.buckets) > peer_mod.NUM_SLOTS - 4:
            raise RuntimeError("too many buckets for the signal pad; raise HVD_FUSION_THRESHOLD")
        cfg = basics.state().config
        forced = cfg.allreduce_variant if cfg else "auto"
        from . import variants as vt
        self.variants, self.ctas = [], []
        for i, b in enumerate(self.buckets):
            v, c = vt.pick(b.nbytes, self.ctx.size, bool(buf.mc_ptr), forced, cfg.comm_ctas if cfg else 0)
            if variants is not None:
                v = variants[i]
            elif cfg is not None and cfg.autotune and forced == "auto" and self.ctx.size > 1 and not self.ctx.emulated:
                v, c = self._autotune(b.nbytes, bool(buf.mc_ptr))
            self.variants.append(v)
            self.ctas.append(c)
        self.side_stream = torch.cuda.Stream(device=arena.device)
        self._events = [torch.cuda.Event() for _ in self.buckets]

    def _autotune(self, nbytes: int, has_nvls: bool):
        """HVD_AUTOTUNE=1: measure every variant x a few CTA counts for this bucket's size on the staging
        buffer (device-timed, max over ranks) and keep the fastest - the measured-latency replacement
        of Horovod's Bayesian autotuner.  Collective: every rank runs the same candidates in order."""
        from .. import collectives
        ctx = self.ctx
        cap = ctx.staging_bytes // 8
        n = min(nbytes // 4, cap) // 4 * 4
        best = None
        log = []
        for v in (ONESHOT, TWOSHOT) + ((NVLS,) if has_nvls and ctx.staging.mc_ptr else ()):
            for ctas in sorted({1, 8, 32, 148}):
                a = HvdFusedArgs()
                a.start, a.end, a.grad_off = 0, n, 0
                a.out_off = cap if v == ONESHOT else 0
                a.scale, a.kind, a.variant, a.slot, a.exit_barrier = 1.0, 0, v, peer_mod.SLOT_STAGING, 1
                for _ in range(2):
                    ctx.launch_fused(ctx.staging, a, ctas)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(5):
                    ctx.launch_fused(ctx.staging, a, ctas)
                e1.record()
                torch.cuda.synchronize()
                t = max(collectives.allgather_object(e0.elapsed_time(e1) / 5))
                log.append({"bytes": n * 4, "variant": VARIANT_NAMES[v], "ctas": ctas, "ms": t})
                if best is None or t < best[0]:
                    best = (t, v, ctas)
        cfg = basics.state().config
        if cfg.autotune_log and ctx.rank == 0:
            import json
            with open(cfg.autotune_log, "a") as f:
                for r in log:
                    f.write(json.dumps(r) + "\n")
        return best[1], best[2]

    def describe(self) -> List[str]:
        return [f"bucket{b.index}[{b.start}:{b.end}] {b.nbytes}B {VARIANT_NAMES[v]} ctas={c}"
                for b, v, c in zip(self.buckets, self.variants, self.ctas)]

    def _args(self, i: int, grad_buf: int, exit_barrier: int, hyper_row=None) -> HvdFusedArgs:
        b, ar = self.buckets[i], self.arena
        a = HvdFusedArgs()
        a.start, a.end = b.start, b.end
        a.grad_off = self.base_off + ar.grad_offset(grad_buf)
        a.param_off = self.base_off + ar.param_offset
        a.slot0_off = self.base_off + (ar.slot_offset(0) if ar.n_slots > 0 else 0)
        a.slot1_off = self.base_off + (ar.slot_offset(1) if ar.n_slots > 1 else 0)
        a.out_off = a.grad_off
        a.hyper = self.opt.hyper_row(hyper_row if hyper_row is not None else 0).data_ptr()
        a.scale = 1.0 / self.ctx.size
        a.kind, a.variant, a.slot, a.exit_barrier = self.opt.kind, self.variants[i], i, exit_barrier
        return a

    def run_bucket(self, i: int, grad_buf: int = 0, exit_barrier: int = 1, stream=None, hyper_row=None) -> None:
        if nvtx.enabled():
            nvtx.push(f"hvd.bucket{i}.{VARIANT_NAMES[self.variants[i]]}")
        self.ctx.launch_fused(self.buf, self._args(i, grad_buf, exit_barrier, hyper_row), self.ctas[i], stream=stream)
        nvtx.pop()

    def run_all(self, grad_buf: int = 0, exit_barrier: int = 1, stream=None) -> None:
        tl = basics.state().timeline
        if tl is None:
            for i in range(len(self.buckets)):
                self.run_bucket(i, grad_buf, exit_barrier, stream)
            return
        # HVD_TIMELINE: device durations from CUDA events around each bucket's kernel ("complete" events,
        # like Horovod's NCCL_ALLREDUCE phase) - costs a sync per step, only when tracing
        s = stream if stream is not None else torch.cuda.current_stream(self.arena.device)
        marks = []
        for i in range(len(self.buckets)):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            self.run_bucket(i, grad_buf, exit_barrier, stream)
            e1.record(s)
            marks.append((i, e0, e1))
        s.synchronize()
        t0 = tl.now_us()
        for i, e0, e1 in marks:
            dur = int(e0.elapsed_time(e1) * 1e3)
            tl.complete(f"grad.bucket{i}", f"FUSED_ALLREDUCE_UPDATE_{VARIANT_NAMES[self.variants[i]].upper()}", t0 - dur, dur)

    def shard_range(self, i: int, rank: int):
        b, n = self.buckets[i], self.ctx.size
        per = (((b.end - b.start) // 4 + n - 1) // n) * 4
        s = min(b.start + rank * per, b.end)
        return s, min(s + per, b.end)

    def materialize_state(self) -> None:
        """Pull the optimizer-slot shards owned by other ranks into this rank's arena (one-sided
        peer reads; called by rank 0 before it writes a checkpoint)."""
        ar = self.arena
        torch.cuda.synchronize(ar.device)
        for i, v in enumerate(self.variants):
            if v == ONESHOT:
                continue
            for r in range(self.ctx.size):
                if r == (0 if self.ctx.emulated else self.ctx.rank):
                    continue
                s, e = self.shard_range(i, r)
                if e <= s:
                    continue
                for k in range(ar.n_slots):
                    o = self.base_off + ar.slot_offset(k)
                    self.ctx.pull(self.buf, r, o + s, e - s)
        torch.cuda.synchronize(ar.device)
