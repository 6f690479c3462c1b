"""Host side of the pipelined data-parallel step (``csrc/comm/pipeline.cuh``, ``pipeline_tail.cu``).

What ``hvd.DistributedOptimizer`` (reference ``mnist_keras.py:87``, ``tensorflow2_keras_mnist.py:58``) turns
into on the fused MNIST engine when the world size divides the 72 row tiles of the Dense(128) kernel:

* the GEMM that produces that kernel's gradient (98% of the bytes) pushes each output tile into the receive
  area of the rank that owns those rows while the conv backward is still running (``tc_gemm`` EPI_RS_PUSH);
* ONE tail kernel per step: a barrier-free LL one-shot allreduce + update of the ~80 KB of small tensors, and the
  owner's reduce + update of its rows (parameters and optimizer state of those rows live with their owner);
* the updated rows reach the other replicas under the NEXT step's conv forward (the first blocks of its conv1
  kernel copy them and stamp version flags; a flush does the same with a stand-alone kernel), and the first
  reader of the big weights waits on those flags - nobody waits in an exit barrier.

This object owns the symmetric receive buffers and fills the argument structs; the engine launches through it.
It also works on an emulated context (``PeerContext.emulate(n)``: n ranks on one GPU, one cooperative launch
of the tail kernel, one GEMM launch per rank) so the kernels can be tested on a single-GPU box.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from ..ops import native
from . import peer as peer_mod
from .peer import MAX_CTAS, MAX_RANKS, HvdCommCtx

SLOT_RS_PUSH, SLOT_TAIL, SLOT_READY = 24, 25, 26
TILE = 128 * 128


class HvdRsPush(ctypes.Structure):
    _fields_ = [("recv", ctypes.c_void_p * MAX_RANKS), ("pad", ctypes.c_void_p * MAX_RANKS),
                ("epoch", ctypes.c_void_p), ("rank", ctypes.c_int), ("nranks", ctypes.c_int),
                ("tiles_per_owner", ctypes.c_int), ("shard_elems", ctypes.c_longlong), ("send_flags", ctypes.c_int)]


class HvdParamWait(ctypes.Structure):
    _fields_ = [("flags", ctypes.c_void_p), ("version", ctypes.c_void_p), ("error", ctypes.c_void_p),
                ("timeout_ns", ctypes.c_ulonglong), ("rank", ctypes.c_int), ("nranks", ctypes.c_int),
                ("ag_ctas", ctypes.c_int), ("trace", ctypes.c_void_p)]


class HvdAllgatherArgs(ctypes.Structure):
    _fields_ = [("w0", ctypes.c_longlong), ("w1", ctypes.c_longlong), ("param_off", ctypes.c_longlong),
                ("version", ctypes.c_void_p * MAX_RANKS), ("use_mc", ctypes.c_int),
                ("trace", ctypes.c_void_p * MAX_RANKS)]


class HvdTailArgs(ctypes.Structure):
    _fields_ = [("a0", ctypes.c_longlong), ("a1", ctypes.c_longlong), ("b0", ctypes.c_longlong),
                ("b1", ctypes.c_longlong), ("w0", ctypes.c_longlong), ("w1", ctypes.c_longlong),
                ("grad_off", ctypes.c_longlong), ("param_off", ctypes.c_longlong),
                ("slot0_off", ctypes.c_longlong), ("slot1_off", ctypes.c_longlong),
                ("recv", ctypes.c_void_p * MAX_RANKS), ("ll", ctypes.c_void_p * MAX_RANKS),
                ("version", ctypes.c_void_p * MAX_RANKS), ("hyper", ctypes.c_void_p),
                ("scale", ctypes.c_float), ("kind", ctypes.c_int), ("parity", ctypes.c_int),
                ("ll_ctas", ctypes.c_int), ("use_mc", ctypes.c_int), ("trace", ctypes.c_void_p * MAX_RANKS)]


def _declare(lib):
    if getattr(lib, "_hvd_pipe_declared", False):
        return
    P = ctypes.POINTER
    lib.hvd_tail_launch.restype = ctypes.c_int
    lib.hvd_tail_launch.argtypes = [P(HvdCommCtx), P(HvdTailArgs), ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    lib.hvd_param_wait_launch.restype = ctypes.c_int
    lib.hvd_param_wait_launch.argtypes = [P(HvdParamWait), ctypes.c_void_p]
    lib.hvd_allgather_launch.restype = ctypes.c_int
    lib.hvd_allgather_launch.argtypes = [P(HvdCommCtx), P(HvdAllgatherArgs), ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_int]
    lib.hvd_rs_flags_launch.restype = ctypes.c_int
    lib.hvd_rs_flags_launch.argtypes = [P(HvdRsPush), ctypes.c_void_p]
    lib._hvd_pipe_declared = True


def supported(nranks: int, big_rows: int) -> bool:
    """The big tensor's 128-row tiles must split evenly over the ranks."""
    return big_rows % 128 == 0 and (big_rows // 128) % nranks == 0 and (big_rows // 128) <= MAX_CTAS


class PipelinedStep:
    """Buffers + argument structs for one model: ``small_a`` / ``small_b`` are the two element ranges of the
    slab that hold the small tensors (in front of and behind the big one), ``big`` is the big tensor's range."""

    def __init__(self, ctx: peer_mod.PeerContext, arena_buf: peer_mod.SymmBuffer, base_off: int, arena, kind: int,
                 small_a, big, small_b, ll_ctas: Optional[int] = None, ctas: Optional[int] = None,
                 ag_ctas: Optional[int] = None):
        self.ctx, self.buf, self.base_off, self.arena, self.kind = ctx, arena_buf, base_off, arena, kind
        self.lib = native.cuda_lib()
        _declare(self.lib)
        n = ctx.size
        self.n = n
        self.small_a, self.big, self.small_b = tuple(small_a), tuple(big), tuple(small_b)
        w = self.big[1] - self.big[0]
        if w % (TILE * n) != 0:
            raise ValueError("the big tensor does not split into whole 128x128 tiles per rank")
        self.shard = w // n
        self.tiles_per_owner = self.shard // TILE
        nsmall = (self.small_a[1] - self.small_a[0]) + (self.small_b[1] - self.small_b[0])
        if nsmall % 4:
            raise ValueError("small ranges must be multiples of 4 elements")
        self.npk = nsmall // 2
        # symmetric receive areas (collective allocations: same order on every rank)
        self.rs = ctx.alloc(w * 4, multicast=False)                      # [src][shard] floats
        self.ll = ctx.alloc(2 * n * self.npk * 16, multicast=False)      # [parity][src][packet] x 16 B
        n_local = n if ctx.emulated else 1
        # number of updates applied to the big tensor (device word per local rank; written by the tail kernel)
        self._versions = [torch.zeros(1, dtype=torch.int32, device=ctx.device) for _ in range(n_local)]
        # all-gather CTAs: ~36 KB of rows each, at most 256 (rank, CTA) delivery flags for the readers to check
        shard_bytes = self.shard * 4
        self.ag_ctas = ag_ctas or max(8, min(64, 256 // max(n, 1), (shard_bytes + 16383) // 16384))
        # One CTA per SM: 256-thread CTAs, the LL part wants one packet per thread, the rest stream the shard
        # LL part: one packet per thread at size > 1 (latency); alone on one GPU it is a trivial local update and
        # the CTAs are better spent on the bandwidth-bound shard part (measured 73.8 vs 74.7 us/step)
        self.ll_ctas = ll_ctas or (8 if n == 1 else max(1, min(64, (self.npk + 255) // 256)))
        total = ctas or 296                # two 256-thread CTAs per SM: the shard part is a bandwidth-bound stream
        if ctx.emulated:
            total = max(2, min(total, (148 * 2) // n))
            self.ll_ctas = max(1, min(self.ll_ctas, total // 2))
        self.ctas = max(total, self.ll_ctas + 1)
        import os
        self.flag_kernel = os.environ.get("HVD_RS_FLAG_KERNEL", "1") != "0"
        self.ag_tma = os.environ.get("HVD_AG_TMA", "1") != "0"      # all-gather by TMA bulk copies (one warp per CTA)
        # multimem.st sends each row once and lets the switch replicate it; with a single peer a plain peer store
        # is the same traffic and measured faster
        self.use_mc = bool(arena_buf.mc_ptr) and n > 1 and os.environ.get("HVD_AG_MC", "1" if n > 2 else "0") != "0"
        self._rs_cache = {}
        self._wait = None
        self._ag_args = None
        # the tail kernel numbers its steps with the context's epoch counter, which may already have advanced
        # (another model trained on this context before): step k of THIS object carries epoch epoch0 + k
        torch.cuda.synchronize(ctx.device)
        self.epoch0 = int(ctx._epochs[0][SLOT_TAIL * MAX_CTAS].item()) & 0xFFFFFFFF
        # device-side trace ring (64 steps x 16 globaltimer stamps) - HVD_TIMELINE on the engine path, profiling
        self._trace = None

    # ------------------------------------------------------------------ structs
    def _rank_index(self, r: Optional[int]) -> int:
        return (r if r is not None else 0) if self.ctx.emulated else 0

    def rs_push(self, rank: Optional[int] = None) -> HvdRsPush:
        """Descriptor for the producing GEMM of ``rank`` (emulation: pass the virtual rank)."""
        key = rank if self.ctx.emulated else -1
        s = self._rs_cache.get(key)
        if s is None:
            s = HvdRsPush()
            for r in range(self.n):
                s.recv[r] = self.rs.ptrs[r]
                s.pad[r] = self.ctx.pad.ptrs[r]
            s.epoch = self.ctx._epochs[self._rank_index(rank)].data_ptr()
            s.rank = rank if self.ctx.emulated else self.ctx.rank
            s.nranks = self.n
            s.tiles_per_owner = self.tiles_per_owner
            s.shard_elems = self.shard
            s.send_flags = 0 if self.flag_kernel else 1
            self._rs_cache[key] = s
        return s

    def launch_rs_flags(self, rank: Optional[int] = None, stream=None) -> None:
        """Publish the tiles of the EPI_RS_PUSH GEMM just enqueued on ``stream`` (no-op when the GEMM's own CTAs
        raise the flags, or at size 1)."""
        if not self.flag_kernel or self.n == 1:
            return
        rc = self.lib.hvd_rs_flags_launch(ctypes.byref(self.rs_push(rank)), self.ctx._stream(stream))
        if rc != 0:
            raise RuntimeError(f"hvd_rs_flags_launch failed: {rc}")

    def flags_ptr(self, rank: Optional[int] = None) -> int:
        r = (rank or 0) if self.ctx.emulated else self.ctx.rank
        return self.ctx.pad.ptrs[r] + 4 * ((SLOT_READY * MAX_CTAS) * MAX_RANKS)

    def param_wait(self, rank: Optional[int] = None) -> Optional[HvdParamWait]:
        """Wait descriptor for kernels that read the big weights (None at size 1: nothing arrives from peers)."""
        if self.n == 1:
            return None
        if self.ctx.emulated or self._wait is None:
            w = HvdParamWait()
            w.flags = self.flags_ptr(rank)
            w.version = self._versions[self._rank_index(rank)].data_ptr()
            w.error = self.ctx._err.data_ptr()
            w.timeout_ns = self.ctx.timeout_ns
            w.rank = (rank or 0) if self.ctx.emulated else self.ctx.rank
            w.nranks, w.ag_ctas = self.n, self.ag_ctas
            w.trace = self._trace[self._rank_index(rank)].data_ptr() if self._trace is not None else None
            if self.ctx.emulated:
                return w
            self._wait = w
        return self._wait

    def tail_args(self, grad_buf: int, parity: int, hyper_ptr: int) -> HvdTailArgs:
        ar, a = self.arena, HvdTailArgs()
        a.a0, a.a1 = self.small_a
        a.b0, a.b1 = self.small_b
        a.w0, a.w1 = self.big
        a.grad_off = self.base_off + ar.grad_offset(grad_buf)
        a.param_off = self.base_off + ar.param_offset
        a.slot0_off = self.base_off + (ar.slot_offset(0) if ar.n_slots > 0 else 0)
        a.slot1_off = self.base_off + (ar.slot_offset(1) if ar.n_slots > 1 else 0)
        for r in range(self.n):
            a.recv[r] = self.rs.ptrs[r]
            a.ll[r] = self.ll.ptrs[r]
            a.version[r] = self._versions[self._rank_index(r)].data_ptr()
            a.trace[r] = self._trace[self._rank_index(r)].data_ptr() if self._trace is not None else None
        a.hyper = hyper_ptr
        a.scale = 1.0 / self.n
        a.kind, a.parity, a.ll_ctas, a.use_mc = self.kind, parity & 1, self.ll_ctas, int(self.use_mc)
        return a

    def enable_trace(self) -> None:
        """Must be called before the step is captured into a CUDA graph (the pointer is a kernel argument).
        The ring lives in page-locked host memory that the kernels write directly (a handful of 8-byte stores per
        step), so the host can read finished records without touching the stream."""
        if self._trace is None:
            self._wait = self._ag_args = None          # cached structs carry the trace pointer
            n_local = self.n if self.ctx.emulated else 1
            self._trace = [torch.zeros(64, 16, dtype=torch.int64).pin_memory() for _ in range(n_local)]
            self._trace_np = [t.numpy() for t in self._trace]

    def read_trace(self, rank: int = 0) -> torch.Tensor:
        """[64][16] int64 nanosecond stamps (rows indexed by step epoch & 63); synchronizes the device."""
        torch.cuda.synchronize(self.ctx.device)
        return self._trace[self._rank_index(rank)].clone()

    def device_clock_offset_us(self, now_us=None) -> float:
        """``now_us()`` (default: the monotonic clock in us) minus the device %globaltimer in us, from one probe
        kernel (error: a launch + sync, ~10 us)."""
        import time
        now_us = now_us or (lambda: time.monotonic() * 1e6)
        probe = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.lib.hvd_globaltimer_probe.restype = ctypes.c_int
        self.lib.hvd_globaltimer_probe.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        rc = self.lib.hvd_globaltimer_probe(probe.data_ptr(), self.ctx._stream(None))
        if rc != 0:
            raise RuntimeError(f"hvd_globaltimer_probe failed: {rc}")
        torch.cuda.synchronize(self.ctx.device)
        return now_us() - float(probe[0]) / 1e3

    # ------------------------------------------------------------------ launches
    def launch_tail(self, grad_buf: int, parity: int, hyper_ptr: int, stream=None) -> None:
        a = self.tail_args(grad_buf, parity, hyper_ptr)
        rc = self.lib.hvd_tail_launch(ctypes.byref(self.ctx.comm_ctx(self.buf)), ctypes.byref(a), self.ctas,
                                      self.ctx._stream(stream), self.n if self.ctx.emulated else 0)
        if rc != 0:
            msg = self.lib.hvd_cuda_error_string(rc).decode() if rc > 0 else f"bad arguments ({rc})"
            raise RuntimeError(f"hvd_tail_launch failed: {msg}")

    def allgather_args(self) -> HvdAllgatherArgs:
        a = self._ag_args
        if a is None:
            a = self._ag_args = HvdAllgatherArgs()
            a.w0, a.w1 = self.big
            a.param_off = self.base_off + self.arena.param_offset
            for r in range(self.n):
                a.version[r] = self._versions[self._rank_index(r)].data_ptr()
                a.trace[r] = self._trace[self._rank_index(r)].data_ptr() if self._trace is not None else None
            a.use_mc = int(self.use_mc)
        return a

    def launch_allgather(self, stream=None) -> None:
        """Deliver this rank's rows of the big tensor (current version) to every other replica.  Idempotent;
        used by ``flush`` (inside a training step the first blocks of the conv1 forward kernel do the same
        through ``allgather_args``).  No-op at size 1."""
        if self.n == 1:
            return
        a = self.allgather_args()
        rc = self.lib.hvd_allgather_launch(ctypes.byref(self.ctx.comm_ctx(self.buf)), ctypes.byref(a), self.ag_ctas,
                                           self.ctx._stream(stream), self.n if self.ctx.emulated else 0,
                                           int(self.ag_tma))
        if rc != 0:
            msg = self.lib.hvd_cuda_error_string(rc).decode() if rc > 0 else f"bad arguments ({rc})"
            raise RuntimeError(f"hvd_allgather_launch failed: {msg}")

    def launch_param_wait(self, stream=None, rank: Optional[int] = None) -> None:
        """Make ``stream`` wait until every peer has delivered the current version of its rows into this arena."""
        w = self.param_wait(rank)
        if w is None:
            return
        rc = self.lib.hvd_param_wait_launch(ctypes.byref(w), self.ctx._stream(stream))
        if rc != 0:
            raise RuntimeError(f"hvd_param_wait_launch failed: {rc}")

    def shard_range(self, rank: int):
        """Element range of the slab whose parameters AND optimizer state rank ``rank`` owns."""
        s = self.big[0] + rank * self.shard
        return s, s + self.shard

    def materialize_state(self) -> None:
        """Pull the optimizer-state rows owned by the other ranks into this rank's arena (one-sided peer
        reads; called before a checkpoint is written).  The caller has already waited for the peers' version flags,
        which every owner stamps only after its tail kernel - and with it its state writes - has completed."""
        ar = self.arena
        me = 0 if self.ctx.emulated else self.ctx.rank
        for r in range(self.n):
            if r == me:
                continue
            s, e = self.shard_range(r)
            for k in range(ar.n_slots):
                self.ctx.pull(self.buf, r, self.base_off + ar.slot_offset(k) + s, e - s)
        torch.cuda.synchronize(ar.device)

    def describe(self):
        w = (self.big[1] - self.big[0]) * 4
        sm = self.npk * 8
        st = "multimem.st" if self.use_mc else "peer stores"
        return [f"big[{self.big[0]}:{self.big[1]}] {w}B: dW GEMM epilogue pushes tiles to their owner (reduce-scatter), "
                f"owner update in the tail (shard ctas={self.ctas - self.ll_ctas}), all-gather by {self.ag_ctas} small CTAs "
                f"under the next step's conv forward ({st}), readers wait on version flags",
                f"small {sm}B: LL one-shot (flagged 16B packets, no barrier), ll ctas={self.ll_ctas}"]
