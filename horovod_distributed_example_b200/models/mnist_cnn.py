"""The fused MNIST-CNN engine: the reference's flagship model (``mnist_keras.py:71-81``,
``tensorflow2_keras_mnist.py:43-52``) as a CUDA-graph-captured chain of hand-written sm_100a
kernels.  Selected automatically by ``Model.compile`` when the layer stack matches (a model that does
not match logs a WARNING and runs on torch autograd).

One training step = ONE graph replay of 9 launches (10 at size > 1), no Python per layer, no host sync:

    conv1_fwd     CUDA cores, K=9; also the step prologue (zeroes the red.add gradient ranges); at size > 1
                  its first blocks deliver last step's updated rows of dense_1/kernel to the other replicas
    conv2_fprop   tcgen05 implicit GEMM + bias/ReLU/maxpool/dropout epilogue; a spare warp of CTA 0 waits for
                  the peers' rows of dense_1/kernel (version flags) so that nothing else has to
    dense1 fwd    tcgen05 split-K GEMM (x[B,9216] * W[9216,128])                              tc_gemm.cu
    head          bias/ReLU/dropout, Dense(10)+softmax+CE+accuracy, dense2 backward, dense1 mask backward
    dense1 dX     tcgen05 GEMM + maxpool/dropout/ReLU-backward scatter epilogue (EPI_UNPOOL)
    dense1 dW     tcgen05 GEMM (both operands MN-major) whose epilogue PUSHES every 128-row tile into the receive
                  area of the rank that owns those rows (EPI_RS_PUSH: GEMM -> reduce-scatter in one kernel,
                  over peer memory), on its own stream under dX and the conv backward; a 1-CTA kernel behind
                  it raises the per-tile flags (size > 1)
    conv2_dgrad   tcgen05 implicit GEMM, zero padding by TMA OOB fill, ReLU-backward epilogue, conv1's
                  weight/bias gradient as a second tensor-core GEMM over tiles staged by the epilogue
    conv2_wgrad   tcgen05, one N=192 MMA per input row covers all nine taps; conv2 bias gradient; CTA 0
                  publishes (loss, accuracy, step) to a page-locked host ring
    tail          csrc/comm/pipeline_tail.cu: barrier-free LL one-shot allreduce + update of the ~80 KB of small
                  tensors, and the owner's reduce + update of its rows of dense_1/kernel (state sharded by owner)

(``HVD_BUCKET_MODE=tail|overlap`` keeps round 1's bucket kernels - ``fused_allreduce.cu`` - selectable; world
sizes that do not divide the 72 row tiles, and ``HVD_DEBUG_POISON``, use them automatically.)

Gradients, inputs and hyper-parameter rows are double-buffered by step parity (two captured graphs).
Loss/accuracy stay on the device: the head kernel writes them into a state block, the wgrad kernel
copies them to the host ring, and the trainer reads them lazily.
"""
from __future__ import annotations

import ctypes
import os
import time
from typing import Optional

import torch

from .. import basics
from ..collectives import _fault_skip, _fault_stall
from ..keras import layers as L
from ..keras import losses as losses_mod
from ..ops import native
from ..utils import nvtx
from ..ops.gemm import EPI_ATOMIC_ADD, EPI_RS_PUSH, EPI_STORE, EPI_UNPOOL, tc_gemm


def _declare(lib):
    if getattr(lib, "_hvd_mnist_declared", False):
        return
    vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    lib.hvd_conv1_fwd.restype = ci
    lib.hvd_conv1_fwd.argtypes = [vp, vp, vp, vp, ci, vp, ci, vp, ci, vp, ci, vp]
    lib.hvd_conv1_fwd_ex.restype = ci
    lib.hvd_conv1_fwd_ex.argtypes = [vp, vp, vp, vp, ci, vp, ci, vp, ci, vp, ci, vp, ci, vp, vp, ci]
    lib.hvd_conv2_fprop_ex.restype = ci
    lib.hvd_conv2_fprop_ex.argtypes = [vp, vp, vp, vp, vp, vp, ci, ci, cf, ci, vp, vp]
    lib.hvd_conv2_fprop.restype = ci
    lib.hvd_conv2_fprop.argtypes = [vp, vp, vp, vp, vp, vp, ci, ci, cf, ci, vp]
    lib.hvd_dense1_fwd_head.restype = ci
    lib.hvd_dense1_fwd_head.argtypes = [vp, vp, ci, ci, ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, cf, vp, vp, vp]
    lib.hvd_head.restype = ci
    lib.hvd_head.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, ci, cf, vp]
    lib.hvd_memcpy_async.restype = ci
    lib.hvd_memcpy_async.argtypes = [vp, vp, ctypes.c_longlong, vp]
    lib.hvd_host_is_pinned.restype = ci
    lib.hvd_host_is_pinned.argtypes = [vp]
    lib.hvd_conv2_dgrad.restype = ci
    lib.hvd_conv2_dgrad.argtypes = [vp, vp, vp, vp, ci, ci, vp, vp, vp, vp]
    lib.hvd_conv2_wgrad.restype = ci
    lib.hvd_conv2_wgrad.argtypes = [vp, vp, vp, vp, ci, ci, vp]
    lib.hvd_conv2_bwd.restype = ci
    lib.hvd_conv2_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, vp, vp, vp, ci, ci]
    lib.hvd_conv2_wgrad_ex.restype = ci
    lib.hvd_conv2_wgrad_ex.argtypes = [vp, vp, vp, vp, ci, ci, vp, vp, vp]
    lib.hvd_conv1_wgrad.restype = ci
    lib.hvd_conv1_wgrad.argtypes = [vp, vp, vp, vp, ci, vp]
    lib.hvd_colsum64.restype = ci
    lib.hvd_colsum64.argtypes = [vp, vp, ctypes.c_longlong, vp]
    lib._hvd_mnist_declared = True


def matches(model) -> bool:
    """True when ``model`` is exactly the reference CNN on 28x28x1 inputs."""
    ls = getattr(model, "layers", None)
    if not ls or len(ls) != 8 or tuple(model.input_shape) != (28, 28, 1):
        return False
    c1, c2, mp, d1, fl, f1, d2, f2 = ls
    if any(getattr(l, "data_format", "channels_last") != "channels_last" for l in ls):
        return False
    ok = (isinstance(c1, L.Conv2D) and c1.filters == 32 and c1.kernel_size == (3, 3) and c1.activation == "relu"
          and c1.padding == "valid" and c1.strides == (1, 1) and c1.use_bias
          and isinstance(c2, L.Conv2D) and c2.filters == 64 and c2.kernel_size == (3, 3) and c2.activation == "relu"
          and c2.padding == "valid" and c2.strides == (1, 1) and c2.use_bias
          and isinstance(mp, L.MaxPooling2D) and mp.pool_size == (2, 2) and mp.strides == (2, 2)
          and isinstance(d1, L.Dropout) and isinstance(fl, L.Flatten)
          and isinstance(f1, L.Dense) and f1.units == 128 and f1.activation == "relu" and f1.use_bias
          and isinstance(d2, L.Dropout)
          and isinstance(f2, L.Dense) and f2.units == 10 and f2.activation == "softmax" and f2.use_bias)
    return bool(ok)


def why_not(model) -> str:
    """Why ``model`` does not run on the fused engine ('' when it does)."""
    if os.environ.get("HVD_ENGINE", "native") != "native":
        return "HVD_ENGINE is not 'native'"
    if model.device.type != "cuda":
        return "not on a CUDA device"
    if not matches(model):
        return "layer stack differs from the reference CNN (Conv32-Conv64-MaxPool-Dropout-Flatten-Dense128-Dropout-Dense10 on 28x28x1)"
    name = losses_mod.loss_name(model.loss) if model.loss is not None else ""
    if "categorical_crossentropy" not in name or getattr(model.loss, "from_logits", False):
        return "loss is not (sparse_)categorical_crossentropy on probabilities"
    if getattr(model.optimizer, "_bpps", 1) != 1:
        return "backward_passes_per_step > 1"
    if torch.cuda.get_device_capability(model.device)[0] < 10:
        return "device is older than sm_100 (tcgen05/TMEM kernels)"
    return ""


def maybe_engine(model):
    """Return a ``MnistEngine`` for ``model`` or None (generic torch-autograd path)."""
    if why_not(model):
        return None
    return MnistEngine(model)


class _StepResult:
    """Handle returned by ``train_host_batch_async``: ``result()`` -> (loss, accuracy) floats.

    A kernel of the step's backward (conv2 wgrad, CTA 0) writes (loss, accuracy, step number) into a
    page-locked host ring; ``result()`` polls that slot's step word - no copy node on the stream, no event."""

    def __init__(self, engine, seq):
        self._e, self._seq, self._val = engine, seq, None

    def ready(self) -> bool:
        return self._val is not None or int(self._e._ring_u32[self._seq & 15][2]) == self._seq

    def result(self):
        if self._val is not None:
            return self._val
        e, seq = self._e, self._seq
        slot_u, slot_f = e._ring_u32[seq & 15], e._ring_np[seq & 15]
        if int(slot_u[2]) != seq:
            if e._train_steps - seq >= 16:
                raise RuntimeError("step result was overwritten: resolve handles within 16 steps")
            t0 = time.monotonic()
            warned = False
            spins = 0
            while int(slot_u[2]) != seq:
                spins += 1
                if spins & 0xfff:
                    continue
                waited = time.monotonic() - t0
                if e.fused is not None:
                    e.fused.ctx.check_error()              # an in-kernel barrier gave up: names the stalled peer
                if not warned and waited > e._stall_warn:
                    warned = True                          # Horovod's stall inspector: a warning, not an abort
                    basics.log.warning("step %d has been waiting for %.0f s (stalled or slow peer?)", seq, waited)
                if waited > e._result_timeout:
                    raise RuntimeError(f"step {seq} did not finish within {e._result_timeout:.0f} s")
        self._val = (float(slot_f[0]), float(slot_f[1]))
        return self._val

    def lazy(self):
        """(loss, accuracy) as ``keras.utils.LazyScalar`` objects for per-batch callback logs."""
        from ..keras.utils import LazyScalar
        return (LazyScalar(lambda: self.result()[0], self.ready), LazyScalar(lambda: self.result()[1], self.ready))


class MnistEngine:
    name = "native-sm100a (tcgen05/TMA kernels, CUDA graph)"

    def __init__(self, model, max_batch: int = 128):
        self.m = model
        self.lib = native.cuda_lib(required=True)
        _declare(self.lib)
        self.dev = model.device
        self.Bmax = max_batch
        ar = model.arena
        c1, c2, _mp, d1, _fl, f1, d2, f2 = model.layers
        self.names = {"c1": c1.name, "c2": c2.name, "f1": f1.name, "f2": f2.name}
        self.keep1, self.keep2 = 1.0 - d1.rate, 1.0 - d2.rate
        B, dev = max_batch, self.dev
        f32 = dict(dtype=torch.float32, device=dev)
        # inputs are double-buffered by step parity (like the gradients): the H2D copy of batch t+1 can
        # land while the graph of step t is still reading batch t
        self.x_bufs = [torch.zeros(B, 28 * 28, **f32) for _ in range(2)]
        self.y_bufs = [torch.zeros(B, dtype=torch.int64, device=dev) for _ in range(2)]
        self.x_in, self.y_in = self.x_bufs[0], self.y_bufs[0]     # parity 0 (also used by eval / predict)
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._dev_index = dev.index if dev.index is not None else torch.cuda.current_device()
        self._main_raw, self._main_obj = None, None
        self._ev_in = [torch.cuda.Event(), torch.cuda.Event()]        # H2D of parity p finished
        self._ev_done = [torch.cuda.Event(), torch.cuda.Event()]      # step that read parity p finished
        self._pin_ring = torch.zeros(16, 4, dtype=torch.float32).pin_memory()      # written by the head kernel
        self._ring_u32 = self._pin_ring.view(torch.int32).numpy().view("uint32")
        self._train_steps = 0                      # mirrors the device step counter (StepState.step)
        # HVD_DEBUG_GUARD=1: every activation buffer sits between two sentinel-filled red zones that
        # check_guards() verifies (compute-sanitizer is not available on every pool; this is the own check)
        self._guards = [] if os.environ.get("HVD_DEBUG_GUARD", "0") == "1" else None
        self.a1 = self._alloc("a1", (B, 26, 26, 32), torch.float32)
        self.p2 = self._alloc("p2", (B, 9216), torch.float32)
        self.mask2 = self._alloc("mask2", (B, 9216), torch.uint8)
        self.h_pre = self._alloc("h_pre", (B, 128), torch.float32).zero_()
        self.dh_pre = self._alloc("dh_pre", (B, 128), torch.float32)
        self.dy2 = self._alloc("dy2", (B, 24, 24, 64), torch.float32)
        self.probs = self._alloc("probs", (B, 10), torch.float32)
        self._pin_x = torch.empty(B, 28 * 28, dtype=torch.float32).pin_memory()
        self._pin_y = torch.empty(B, dtype=torch.int64).pin_memory()
        self._pin_x_np, self._pin_y_np = self._pin_x.numpy(), self._pin_y.numpy()
        self._ring_np = self._pin_ring.numpy()
        self._ring_ptr = self._pin_ring.data_ptr()
        self._x_ptr = [t.data_ptr() for t in self.x_bufs]
        self._y_ptr = [t.data_ptr() for t in self.y_bufs]
        self._pin_cache = {}
        # StepState: step, loss_sum, correct_sum, last_loss, last_correct, seed, ring idx pads
        self.state = torch.zeros(8, dtype=torch.int32, device=dev)
        seed = (model._seed if model._seed is not None else int.from_bytes(os.urandom(3), "little")) & 0x7FFFFFFF
        rank = basics.rank() if basics.is_initialized() else 0
        self.state[5] = (seed * 2654435761 + 97 * rank + 1) & 0x7FFFFFFF     # per-rank dropout masks
        self.state_f = self.state.view(torch.float32)
        self._graphs = {}
        self._parity = 0
        self._ag_pending = False
        self._conv1_bps = 8
        self._tl_offset, self._tl_done, self._tl_active = None, 0, False
        self._hyper_pushed = False
        # HVD_FUSE_HEAD=1: dense1 forward + head as one launch (EPI_SPLITK_HEAD)
        self._fuse_head = os.environ.get("HVD_FUSE_HEAD", "0") == "1"
        self._head_ctr = torch.zeros(1, dtype=torch.int64, device=dev)
        # where the all-gather of the updated rows runs: "conv1" = as the first blocks of the next step's conv1
        # forward kernel (default), "start" = a separate small kernel forked at the start of the next step,
        # "tail" = right behind the tail kernel (exposed; A/B only)
        self._ag_where = os.environ.get("HVD_AG_WHERE", "conv1")
        self._wait_in = os.environ.get("HVD_WAIT_IN", "fprop")
        self.side = torch.cuda.Stream(device=dev)
        self.side2 = torch.cuda.Stream(device=dev)
        self._ev_fork = torch.cuda.Event()
        self._ev_dx = torch.cuda.Event()
        self._ev_b = torch.cuda.Event()
        opt = model.optimizer
        self.fused = getattr(opt, "_fused", None)
        # HVD_BUCKET_MODE=tail: ONE bucket reduced + updated after the whole backward.
        # =overlap: bucket 0 = dense layers (98% of the bytes) goes to a side stream as soon as dense1's dW is
        # done, while the conv backward runs.  The conv backward CTAs fill an SM's registers and shared memory,
        # so the overlap only works if some SMs are left free for the bucket's CTAs: HVD_RESERVE_SMS (default
        # 16) SMs are kept out of the dgrad/wgrad split and bucket 0 runs with two CTAs on each of them.
        # Measured: tail wins at 1 GPU (79.6 vs 87.7 us/step without a reservation) and at 2 GPUs (110 us/step
        # vs 134 / 115 with 16 / 24 SMs reserved), so tail is the default; overlap stays selectable.
        # HVD_BUCKET_MODE=pipelined (default): no bucket kernels at all.  dense1's dW GEMM pushes its tiles to
        # their owner ranks while the conv backward runs, ONE tail kernel does the barrier-free LL allreduce +
        # update of the small tensors and the owners' reduce + update of the big one, the next step's conv1 kernel
        # delivers the updated rows to the other replicas, and that
        # step's conv2 kernel waits for the delivery flags instead of anybody sitting in an exit barrier
        # (parallel/pipeline.py).  Needs the 72 row tiles of dense_1/kernel to split evenly over the ranks.
        from ..parallel import pipeline as pipe_mod
        mode = os.environ.get("HVD_BUCKET_MODE", "pipelined")
        if mode not in ("tail", "overlap", "pipelined"):
            mode = "pipelined"
        if mode == "pipelined" and (self.fused is None or not pipe_mod.supported(self.fused.ctx.size, 9216)
                                    or self._debug_modes()):
            mode = "tail"
        self._bucket_mode = mode
        self.pipe = None
        reserve = int(os.environ.get("HVD_RESERVE_SMS", 16)) if self._bucket_mode == "overlap" else 0
        d_def, w_def = (110, 38) if reserve == 0 else (max(148 - reserve - 32, 8), 32)
        self._dgrad_ctas = int(os.environ.get("HVD_DGRAD_CTAS", d_def))     # dgrad and wgrad run side by side: split the SMs
        self._wgrad_ctas = int(os.environ.get("HVD_WGRAD_CTAS", w_def))
        # HVD_FUSED_BWD=1: conv2 wgrad + dgrad as ONE kernel (hvd_conv2_bwd).  Measured equal to the two kernels
        # side by side (75.8 us/step both; the pair is bound by L2/TMA latency, not by the SM split), so the
        # default stays the two-kernel form that the multi-GPU runs validated.
        self._fused_bwd = os.environ.get("HVD_FUSED_BWD", "0") == "1" and reserve == 0
        self._bwd_ctas = int(os.environ.get("HVD_BWD_CTAS", 148))
        self._bwd_nw = int(os.environ.get("HVD_BWD_NW", 0))          # CTAs that reduce wgrad strips first (0 = default)
        self._bwd_t1 = int(os.environ.get("HVD_BWD_T1", -1))         # dgrad tiles those CTAs still take (-1 = default)
        if hasattr(opt, "set_bucket_cuts") and self.fused is not None:
            opt.set_bucket_cuts([f"{self.names['f1']}/kernel"] if self._bucket_mode == "overlap" else [])
            self.fused = opt._fused
            if reserve > 0 and not basics.state().config.comm_ctas:
                self.fused.ctas[0] = min(self.fused.ctas[0], 2 * reserve)
        if self._bucket_mode == "pipelined":
            ar = model.arena
            big = ar.spec(f"{self.names['f1']}/kernel")
            w0, w1 = big.offset, big.offset + big.numel
            self.pipe = pipe_mod.PipelinedStep(self.fused.ctx, self.fused.buf, self.fused.base_off, ar, opt.kind,
                                               (0, w0), (w0, w1), (w1, ar.n),
                                               ll_ctas=int(os.environ.get("HVD_LL_CTAS", 0)) or None,
                                               ctas=int(os.environ.get("HVD_TAIL_CTAS", 0)) or None,
                                               ag_ctas=int(os.environ.get("HVD_AG_CTAS", 0)) or None)
            opt._pipe = self.pipe             # checkpoints gather the owner-sharded state through it
            self._conv1_bps = int(os.environ.get("HVD_CONV1_BPS", 8))
            if os.environ.get("HVD_TAIL_TRACE", "0") == "1" or basics.state().timeline is not None:
                self.pipe.enable_trace()
            if basics.state().timeline is not None:
                self._tl_active = True
                basics.state().shutdown_hooks.append(lambda: self.drain_timeline(final=True))
        elif hasattr(opt, "_pipe"):
            opt._pipe = None
        cfg = basics.state().config
        self._fault = cfg.parsed_fault() if cfg is not None else None
        self._debug = cfg is not None and bool(cfg.debug_checksum or cfg.debug_poison)
        self._stall_warn = float(cfg.stall_check_time) if cfg is not None and cfg.stall_check_time > 0 else 60.0
        abort = cfg.barrier_abort_time() if cfg is not None else 0.0
        self._result_timeout = (abort + 5.0) if abort > 0 else 3600.0     # the in-kernel abort fires first
        self.use_graph = os.environ.get("HVD_NO_GRAPH", "0") != "1" and self.fused is not None
        if self.fused is not None:
            opt._defer_push = True           # hyper-parameter rows are written at step start (see _run_step)
        if self.pipe is not None:      # tail kernel (+ the tile-flag kernel behind dense1's dW at size > 1)
            comm_launches = 1 + (1 if (self.pipe.flag_kernel and self.pipe.n > 1) else 0)
        else:
            comm_launches = len(self.fused.buckets) if self.fused is not None else 0
        self.launches_per_step = (7 if self._fused_bwd else 8) - (1 if self._fuse_head else 0) + comm_launches

    # ------------------------------------------------------------------ helpers
    @staticmethod
    def _debug_modes() -> bool:
        """HVD_DEBUG_POISON NaN-fills the local gradient slab after each update, which the pipelined step's
        big gradient never visits: those runs use the bucket kernels."""
        cfg = basics.state().config
        return cfg is not None and bool(cfg.debug_poison)

    def describe_comm(self):
        if self.pipe is not None:
            return self.pipe.describe()
        return self.fused.describe() if self.fused is not None else ["single process, unfused optimizer"]

    _GUARD_BYTES = 4096

    def _alloc(self, name, shape, dtype):
        n = 1
        for d in shape:
            n *= d
        if self._guards is None:
            return torch.empty(shape, dtype=dtype, device=self.dev)
        esz = torch.empty(0, dtype=dtype).element_size()
        gb = self._GUARD_BYTES
        raw = torch.full((2 * gb + n * esz,), 0xA5, dtype=torch.uint8, device=self.dev)
        self._guards.append((name, raw, gb, n * esz))
        return raw[gb:gb + n * esz].view(dtype).view(shape)

    def check_guards(self):
        """Raise if any kernel wrote outside its activation buffer (HVD_DEBUG_GUARD=1 only)."""
        if self._guards is None:
            raise RuntimeError("guards are off: set HVD_DEBUG_GUARD=1 before the model is compiled")
        torch.cuda.synchronize(self.dev)
        for name, raw, gb, nb in self._guards:
            for side, zone in (("below", raw[:gb]), ("above", raw[gb + nb:])):
                bad = (zone != 0xA5).nonzero()
                if bad.numel():
                    raise RuntimeError(f"out-of-bounds write {side} buffer '{name}': {bad.numel()} byte(s), "
                                       f"first at offset {int(bad[0])} of the red zone")
        return True

    def _chk(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed with code {rc}")

    _SKIP = set(filter(None, os.environ.get("HVD_ABLATE", "").split(",")))   # profiling only: drop kernels from the step

    def _s(self):
        return torch._C._cuda_getCurrentRawStream(self._dev_index)

    def _main(self):
        """The current stream as a Stream object.  ``torch.cuda.current_stream`` costs several microseconds of
        Python per call and the per-step path needs it twice, so the object is cached and only re-created
        when the raw handle of the current stream has changed."""
        raw = torch._C._cuda_getCurrentRawStream(self._dev_index)
        if raw != self._main_raw:
            self._main_obj = torch.cuda.current_stream(self.dev)
            self._main_raw = raw
        return self._main_obj

    def _p(self, key, short):
        return self.m.arena.param(f"{self.names[key]}/{short}")

    def _g(self, key, short, which):
        return self.m.arena.grad(f"{self.names[key]}/{short}", which)

    def weights_changed(self):
        pass                                  # kernels read the arena directly; nothing is cached

    # ------------------------------------------------------------------ forward
    def _forward(self, B: int, training: bool, which: Optional[int] = None):
        self._forward_conv(B, training, which)
        self._forward_dense(B)

    def _forward_dense(self, B: int):
        if "d1fwd" not in self._SKIP:
            # pipelined mode: the first reader of dense_1/kernel waits until its owners' stores have landed
            tc_gemm(self.p2, self._p("f1", "kernel"), B, 128, 9216, a_mn=False, b_mn=True, bn=128,
                    k_split=int(os.environ.get('HVD_D1_KSPLIT', 32)), epilogue=EPI_ATOMIC_ADD, out=self.h_pre,
                    stream=self._s(),
                    wait=self.pipe.param_wait() if (self.pipe is not None and self._wait_in == "d1fwd") else None)

    def _forward_conv(self, B: int, training: bool, which: Optional[int] = None):
        lib, s = self.lib, self._s()
        za = zb = None
        na = nb = 0
        if which is not None:
            # ranges accumulated with red.add this step: the dense2/bias1 grads in front of dense_1/kernel
            # and the conv grads behind it (the arena is in backward-ready order)
            ar = self.m.arena
            g = ar.grad_bufs[which]
            na = ar.spec(f"{self.names['f1']}/kernel").offset
            first_conv = ar.spec(f"{self.names['c2']}/bias").offset
            nb = ar.n - first_conv
            za, zb = g.data_ptr(), g.data_ptr() + 4 * first_conv
        xin = self.x_bufs[which or 0]
        if "conv1_fwd" not in self._SKIP:
            cc = ag = None
            agb = 0
            if which is not None and self.pipe is not None and self.pipe.n > 1 and self._ag_where == "conv1":
                # the first blocks of this kernel deliver last step's updated rows of dense_1/kernel to the peers
                cc = ctypes.addressof(self.pipe.ctx.comm_ctx(self.pipe.buf))
                ag = ctypes.addressof(self.pipe.allgather_args())
                agb = self.pipe.ag_ctas
            self._chk(lib.hvd_conv1_fwd_ex(xin.data_ptr(), self._p("c1", "kernel").data_ptr(),
                                           self._p("c1", "bias").data_ptr(), self.a1.data_ptr(), B, za, na, zb, nb,
                                           self.state.data_ptr(), int(which is not None), s,
                                           self._conv1_bps if which is not None else 8, cc, ag, agb), "conv1_fwd")
        if "fprop" not in self._SKIP:
            # pipelined mode: CTA 0's spare warp also waits for the peers' rows of dense_1/kernel, so the Dense
            # forward behind this kernel starts with the weights already delivered (HVD_WAIT_IN=d1fwd moves it back)
            w = self.pipe.param_wait() if (self.pipe is not None and self._wait_in == "fprop") else None
            self._chk(lib.hvd_conv2_fprop_ex(self.a1.data_ptr(), self._p("c2", "kernel").data_ptr(),
                                             self._p("c2", "bias").data_ptr(), self.p2.data_ptr(), self.mask2.data_ptr(),
                                             self.state.data_ptr(), B, int(training), self.keep1, 0, s,
                                             ctypes.addressof(w) if w is not None else None), "conv2_fprop")

    def _head(self, B: int, training: bool, backward: bool, which: int = 0, probs: bool = False):
        gp = (lambda k, sh: self._g(k, sh, which).data_ptr()) if backward else (lambda k, sh: None)
        if "head" in self._SKIP:
            return
        labels = self.y_bufs[which if backward else 0]
        self._chk(self.lib.hvd_head(self.h_pre.data_ptr(), self._p("f1", "bias").data_ptr(),
                                    self._p("f2", "kernel").data_ptr(), self._p("f2", "bias").data_ptr(),
                                    labels.data_ptr(), self.dh_pre.data_ptr(), gp("f2", "kernel"),
                                    gp("f2", "bias"), gp("f1", "bias"),
                                    self.probs.data_ptr() if probs else None, self.state.data_ptr(), B,
                                    int(training), int(backward), self.keep2, self._s()), "head")

    # ------------------------------------------------------------------ the training step
    # One step = one captured graph with three parts:
    #   A: conv1_fwd, conv2_fprop
    #   B: dense1 fwd, head, then dense1 dW || dense1 dX
    #   [bucket 0: fused allreduce+update of the dense layers (98% of the bytes), forked onto the side stream]
    #   C: conv2_dgrad(+conv1 wgrad) || conv2_wgrad(+bias grad), then bucket 1 (conv layers); join.
    # (Splitting the step into three graphs so that bucket 0 also overlaps the NEXT step's part A was
    # measured and lost: 148 vs 136 us/step at 2 GPUs - three replays plus the out-of-graph launch made
    # the host the bottleneck.  Gradients and hyper-parameters stay double-buffered by step parity.)
    def _part_a(self, B: int, which: int):
        self._forward_conv(B, True, which)

    def _dense_head_fused(self, B: int, which: int):
        """dense1 forward (split-K) + the head in ONE launch (tc_gemm EPI_SPLITK_HEAD): the slice CTAs wait for each
        other on a ticket counter and the first 16 of them run the head on 8 rows each."""
        gp = lambda k, sh: self._g(k, sh, which).data_ptr()                              # noqa: E731
        w = self.pipe.param_wait() if (self.pipe is not None and self._wait_in == "d1fwd") else None
        self._chk(self.lib.hvd_dense1_fwd_head(
            self.p2.data_ptr(), self._p("f1", "kernel").data_ptr(), B, 9216, int(os.environ.get('HVD_D1_KSPLIT', 32)),
            self.h_pre.data_ptr(), self._p("f1", "bias").data_ptr(), self._p("f2", "kernel").data_ptr(),
            self._p("f2", "bias").data_ptr(), self.y_bufs[which].data_ptr(), self.dh_pre.data_ptr(), gp("f2", "kernel"),
            gp("f2", "bias"), gp("f1", "bias"), None, self.state.data_ptr(), 1, 1, self.keep2,
            self._head_ctr.data_ptr(), ctypes.addressof(w) if w is not None else None, self._s()), "dense1_fwd_head")

    def _part_b(self, B: int, which: int):
        main = torch.cuda.current_stream(self.dev)
        if self._fuse_head and not ({"d1fwd", "head"} & self._SKIP):
            self._dense_head_fused(B, which)
        else:
            self._forward_dense(B)
            self._head(B, True, True, which)
        self._ev_fork.record(main)
        if self.pipe is not None:
            # GEMM -> reduce-scatter in one kernel: every 128-row tile goes to its owner's receive area (local
            # memory or a peer over NVLink).  On its own stream and joined only before the tail kernel, so its
            # peer stores + system fence drain under dX and the conv backward instead of delaying them.
            self.side.wait_event(self._ev_fork)
            if "dw" not in self._SKIP:
                with torch.cuda.stream(self.side):
                    tc_gemm(self.p2, self.dh_pre, 9216, 128, B, a_mn=True, b_mn=True, bn=128, epilogue=EPI_RS_PUSH,
                            out=None, stream=self.side.cuda_stream, rs=self.pipe.rs_push())
                    self.pipe.launch_rs_flags(stream=self.side)
            if "dx" not in self._SKIP:
                tc_gemm(self.dh_pre, self._p("f1", "kernel"), B, 9216, 128, a_mn=False, b_mn=False, bn=64,
                        epilogue=EPI_UNPOOL, out=self.dy2.view(-1, 64), mask=self.mask2, scale=1.0 / self.keep1,
                        stream=self._s())
            return
        self.side2.wait_event(self._ev_fork)
        with torch.cuda.stream(self.side2):
            if "dw" not in self._SKIP:
                tc_gemm(self.p2, self.dh_pre, 9216, 128, B, a_mn=True, b_mn=True, bn=128, epilogue=EPI_STORE,
                        out=self._g("f1", "kernel", which), stream=self.side2.cuda_stream)
        if "dx" not in self._SKIP:
            tc_gemm(self.dh_pre, self._p("f1", "kernel"), B, 9216, 128, a_mn=False, b_mn=False, bn=64,
                    epilogue=EPI_UNPOOL, out=self.dy2.view(-1, 64), mask=self.mask2, scale=1.0 / self.keep1,
                    stream=self._s())
        main.wait_stream(self.side2)

    def _part_c(self, B: int, which: int):
        lib, fused = self.lib, self.fused
        main = torch.cuda.current_stream(self.dev)
        s = self._s()
        fused_one = self._fused_bwd and not ({"dgrad", "wgrad"} & self._SKIP)
        if fused_one:
            # ONE kernel: every CTA does its share of conv2 wgrad strips, then its share of dgrad tiles (with
            # conv1's ReLU backward and weight/bias gradient fused in); CTA 0 also publishes the step's
            # loss/accuracy into the pinned host ring.  dense1's dW (side2) only has to be done before the update.
            self._chk(lib.hvd_conv2_bwd(self.dy2.data_ptr(), self._p("c2", "kernel").data_ptr(), self.a1.data_ptr(),
                                        self.x_bufs[which].data_ptr(), self._g("c2", "kernel", which).data_ptr(),
                                        self._g("c2", "bias", which).data_ptr(), self._g("c1", "kernel", which).data_ptr(),
                                        self._g("c1", "bias", which).data_ptr(), B, self._bwd_ctas, s,
                                        self.state.data_ptr(), self._ring_ptr, self._bwd_nw, self._bwd_t1), "conv2_bwd")
        else:
            self._ev_dx.record(main)
            self.side2.wait_event(self._ev_dx)
            with torch.cuda.stream(self.side2):
                if "wgrad" not in self._SKIP:
                    # (CTA 0 of this kernel also writes the step's loss/accuracy into the pinned host ring)
                    self._chk(lib.hvd_conv2_wgrad_ex(self.a1.data_ptr(), self.dy2.data_ptr(),
                                                     self._g("c2", "kernel", which).data_ptr(),
                                                     self._g("c2", "bias", which).data_ptr(), B, self._wgrad_ctas,
                                                     self.side2.cuda_stream, self.state.data_ptr(), self._ring_ptr), "conv2_wgrad")
            # conv2 dgrad with conv1's ReLU backward AND conv1's weight/bias gradient fused in (da1 never
            # touches global memory)
            if "dgrad" not in self._SKIP:
                self._chk(lib.hvd_conv2_dgrad(self.dy2.data_ptr(), self._p("c2", "kernel").data_ptr(), self.a1.data_ptr(),
                                              None, B, self._dgrad_ctas, s, self.x_bufs[which].data_ptr(),
                                              self._g("c1", "kernel", which).data_ptr(),
                                              self._g("c1", "bias", which).data_ptr()), "conv2_dgrad")
        if not (fused_one and self.pipe is not None):
            main.wait_stream(self.side2)         # (pipelined + one-kernel conv backward: nothing was forked onto side2)
        if self.pipe is not None:
            main.wait_stream(self.side)          # dense1 dW (pushed to its owners) joins here
            if "b1" not in self._SKIP:
                self.pipe.launch_tail(which, which, self.m.optimizer.hyper_row(which).data_ptr())
                if self._ag_where == "tail":     # A/B switch: deliver the rows right behind the tail (exposed)
                    self.pipe.launch_allgather()
        elif fused is not None and "b1" not in self._SKIP:
            for i in range(1 if self._bucket_mode == "overlap" else 0, len(fused.buckets)):
                fused.run_bucket(i, which, self._exit_barrier(), hyper_row=which)

    def _exit_barrier(self) -> int:
        return 0 if self.m.arena.n_grad_bufs == 2 else 1

    def sync_params(self):
        """Make the current stream wait for any update still running on the side stream (call before
        anything outside the training step reads parameters or optimizer state)."""
        torch.cuda.current_stream(self.dev).wait_stream(self.side)
        if self.pipe is not None:
            self.flush_params()
            # ... and wait for the rows that the OTHER ranks deliver into this arena (their next training step,
            # their own sync or their end-of-epoch flush sends them)
            self.pipe.launch_param_wait()

    # ------------------------------------------------------------------ HVD_TIMELINE on the engine path
    _TL_EVENTS = ((0, 3, "grad.small_tensors", "FUSED_ALLREDUCE_UPDATE_LL"),
                  (4, 7, "grad.dense_1/kernel", "FUSED_ALLREDUCE_UPDATE_OWNER_ROWS"),
                  (10, 11, "param.dense_1/kernel", "ALLGATHER_UPDATED_ROWS"),
                  (12, 13, "param.dense_1/kernel", "WAIT_PEER_ROWS"))

    def drain_timeline(self, final: bool = False):
        """Write the tail kernel's own device timestamps (pipeline_tail.cu: %globaltimer stamps in a page-locked
        ring) into the Chrome trace as complete events, one set per training step.  Lazy and sync-free: only
        records at least 12 steps behind the launches are read (``final`` synchronizes and takes the rest)."""
        tl = basics.state().timeline
        if tl is None or self.pipe is None or self.pipe._trace is None:
            return
        if self._tl_offset is None:
            self._tl_offset = self.pipe.device_clock_offset_us(tl.now_us)
        if final:
            torch.cuda.synchronize(self.dev)
        ring = self.pipe._trace_np[0]
        upto = self._train_steps if final else self._train_steps - 12
        while self._tl_done < upto:
            step = self._tl_done + 1
            if self._train_steps - step >= 60:         # overwritten by newer steps before it was read
                self._tl_done += 1
                continue
            ep = (self.pipe.epoch0 + step) & 0xFFFFFFFF
            row = ring[ep & 63]
            if int(row[15]) != ep:
                break
            for a, b, tensor, name in self._TL_EVENTS:
                t0, t1 = int(row[a]), int(row[b])
                # the all-gather / wait stamps of the LAST step only exist once another step (or a flush) ran
                if t0 > 0 and t1 >= t0 and (a < 8 or t0 >= int(row[0])):
                    tl.complete(tensor, name, int(t0 / 1e3 + self._tl_offset), max(1, int((t1 - t0) / 1e3)))
            self._tl_done += 1

    def flush_params(self):
        """Pipelined mode: deliver the rows of dense_1/kernel this rank owns to the other replicas if a training
        step has updated them since the last delivery (normally the NEXT step does it, hidden under its conv
        forward).  Non-blocking, one small launch; ``fit`` calls it at every epoch end and the blocking
        ``train_on_batch`` after every step, so that a rank reading weights never waits on a peer that has
        simply stopped training."""
        if self.pipe is not None and self._ag_pending:
            self.pipe.launch_allgather()
            self._ag_pending = False

    def _stage_host(self, x, y):
        """Enqueue the H2D copy of one host batch into the input buffers of the NEXT step's parity on the
        copy stream (straight from the caller's memory when it is page-locked, else through a pinned staging
        pair) and make the main stream wait for it.  Returns the batch size.

        This runs once per step next to an ~80 us device step, so it avoids framework tensors: two
        ``cudaMemcpyAsync`` through the C ABI on raw pointers, and the page-locked test is cached per owning
        buffer (a wrong cache hit only makes the copy synchronous, never incorrect)."""
        import numpy as np
        xa = x if (x.dtype == np.float32 and x.flags.c_contiguous) else np.ascontiguousarray(x, dtype=np.float32)
        if y.ndim == 2:
            y = y.argmax(axis=1)
        ya = y if (y.dtype == np.int64 and y.flags.c_contiguous) else np.ascontiguousarray(y, dtype=np.int64)
        B = xa.shape[0]
        if B > self.Bmax:
            raise ValueError(f"batch {B} exceeds the engine's max batch {self.Bmax}")
        if xa.size != B * 784:
            raise ValueError(f"expected {B} x 28 x 28 x 1 inputs, got shape {tuple(xa.shape)}")
        w = self._which()
        main = self._main()
        cs = self.copy_stream
        xp, yp = xa.ctypes.data, ya.ctypes.data
        if not (self._pinned(xa, xp) and self._pinned(ya, yp)):
            cs.synchronize()                      # the pinned staging pair is about to be overwritten
            self._pin_x_np[:B] = xa.reshape(B, 784)
            self._pin_y_np[:B] = ya
            xp, yp = self._pin_x.data_ptr(), self._pin_y.data_ptr()
        cs.wait_event(self._ev_done[w])           # the step that last read this parity's buffers is finished
        csh = cs.cuda_stream
        lib = self.lib
        rc = lib.hvd_memcpy_async(self._x_ptr[w], xp, B * 3136, csh) or lib.hvd_memcpy_async(self._y_ptr[w], yp, B * 8, csh)
        if rc:
            raise RuntimeError(f"input staging copy failed with CUDA error {rc}")
        opt = self.m.optimizer
        if self.fused is not None and (opt._hyper_dirty > 0 or opt.kind == 3 or getattr(opt, "decay", 0.0) > 0.0):
            # this step's row of the hyper-parameter block travels with the batch on the copy stream (the row's
            # last readers were step t-2's update kernels, which _ev_done[w] above has already covered)
            opt._push_hyper_raw(w, lib.hvd_memcpy_async, cs)
            opt._hyper_dirty = max(0, opt._hyper_dirty - 1)
            self._hyper_pushed = True
        self._ev_in[w].record(cs)
        main.wait_event(self._ev_in[w])
        return B

    def _pinned(self, arr, ptr) -> bool:
        owner = arr.base if arr.base is not None else arr
        key = id(owner)
        hit = self._pin_cache.get(key)
        if hit is None:
            if len(self._pin_cache) > 64:
                self._pin_cache.clear()
            hit = self._pin_cache[key] = bool(self.lib.hvd_host_is_pinned(ptr))
        return hit

    def _which(self) -> int:
        return self._parity if self.m.arena.n_grad_bufs == 2 else 0

    def train_host_batch_async(self, x, y):
        """Pipelined end-to-end step from HOST arrays: H2D of this batch (copy stream), the step's graph,
        which writes (loss, accuracy, step) into a pinned host ring - all enqueued without blocking.  Returns a
        handle; ``handle.result()`` waits for THAT step only.  Calling it one step late lets the H2D copy
        of batch t+1 overlap the compute of step t."""
        B = self._stage_host(x, y)
        self._run_step(B)
        return _StepResult(self, self._train_steps)

    def train_host_lazy(self, x, y):
        """fit()'s path: pipelined H2D + step, returning DEVICE scalars (no host sync at all)."""
        B = self._stage_host(x, y)
        self._run_step(B)
        return self.state_f[3], self.state_f[4]

    def train_host_batch(self, x, y):
        """Blocking end-to-end step from HOST arrays; returns (loss, accuracy) as Python floats."""
        out = self.train_host_batch_async(x, y).result()
        self.flush_params()
        return out

    def train_step(self, x: torch.Tensor, y: torch.Tensor):
        B = x.shape[0]
        if B > self.Bmax:
            raise ValueError(f"batch {B} exceeds the engine's max batch {self.Bmax}")
        if y.ndim == 2:
            y = y.argmax(dim=1)
        w = self._which()
        self.x_bufs[w][:B].copy_(x.reshape(B, -1), non_blocking=True)
        self.y_bufs[w][:B].copy_(y.to(torch.int64), non_blocking=True)
        self._run_step(B)
        # Device scalars, read lazily by the trainer (no host sync, no extra launch).  They are views
        # of the engine's state block: valid until the next step overwrites them - consumers (the fit
        # loop's device-side running sums, callbacks calling float()) use them immediately.
        return self.state_f[3], self.state_f[4]

    def _run_step(self, B: int):
        opt = self.m.optimizer
        fused = self.fused
        if hasattr(opt, "begin_step"):
            opt.begin_step()
        if fused is not None:
            fused.ctx.check_error()          # host watchdog (reads one pinned word; no sync)
            if self._fault is not None and _fault_skip():   # HVD_FAULT_INJECT=rank:step: this rank hangs, then fails
                _fault_stall()
        which = self._which()
        main = self._main()
        nvtx.push("hvd.step")
        if self._hyper_pushed:
            self._hyper_pushed = False           # _stage_host already sent this step's row on the copy stream
        elif fused is not None and (opt._hyper_dirty > 0 or opt.kind == 3 or getattr(opt, "decay", 0.0) > 0.0):
            # this step's row of the hyper-parameter block (LR changed by a callback, Adam's lr_t, LR decay);
            # nothing in flight reads this row: its last readers were step t-2's update kernels
            opt._next_slot = which
            opt._push_hyper()
            opt._hyper_dirty = max(0, opt._hyper_dirty - 1)
        if self.use_graph:
            g = self._graphs.get((B, which))
            if g is None:
                g = self._graphs[(B, which)] = self._capture(B, which)
            g.replay()
        else:
            self._whole_step(B, which)
            if fused is None:
                if which != 0:
                    raise RuntimeError("unfused optimizer path expects a single gradient buffer")
                opt.apply_gradients(self.m.arena)
        self._ev_done[which].record(main)
        self._ag_pending = True
        if fused is not None:
            opt.iterations += 1
            self._parity ^= 1 if self.m.arena.n_grad_bufs == 2 else 0
            if self._debug:
                self.sync_params()           # the replica checksum must see the weights peers are still sending
                opt._debug_after_step(self.m.arena, which)
        self._train_steps += 1
        if self._tl_active and (self._train_steps & 15) == 0:
            self.drain_timeline()
        nvtx.pop()

    def _whole_step(self, B: int, which: int):
        """Enqueue the whole step on the current stream (capturable): A, B, bucket 0 forked onto the side
        stream, C, join."""
        main = torch.cuda.current_stream(self.dev)
        fused = self.fused
        if self.pipe is not None and self.pipe.n > 1 and self._ag_where == "start":
            # the rows of dense_1/kernel this rank updated in the previous step travel to the other replicas NOW,
            # from a few small CTAs that share the SMs with the conv forward; dense1's forward waits on their flags
            self._ev_b.record(main)
            self.side.wait_event(self._ev_b)
            self.pipe.launch_allgather(stream=self.side)
        self._part_a(B, which)
        self._part_b(B, which)
        forked = False
        if fused is not None and "b0" not in self._SKIP and self._bucket_mode == "overlap":
            # bucket 0 (dense layers, 98% of the bytes) is complete: reduce + update it on the side stream
            # while the conv backward runs.  It rewrites dense_1/kernel, which dX (end of part B) has
            # finished reading by now.
            self._ev_b.record(main)
            self.side.wait_event(self._ev_b)
            fused.run_bucket(0, which, self._exit_barrier(), stream=self.side, hyper_row=which)
            forked = True
        self._part_c(B, which)
        if forked:
            main.wait_stream(self.side)

    def _capture(self, B: int, which: int):
        import gc
        gc.collect()                      # a stale CUDAGraph finalised mid-capture would invalidate the capture
        torch.cuda.synchronize(self.dev)
        g = torch.cuda.CUDAGraph()
        cap = torch.cuda.Stream(device=self.dev)
        cap.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.graph(g, stream=cap, capture_error_mode="thread_local"):
            self._whole_step(B, which)
        torch.cuda.current_stream(self.dev).wait_stream(cap)
        return g

    # ------------------------------------------------------------------ eval / predict
    def _inference(self, B: int, probs: bool):
        """Forward + head with dropout off, replayed from a CUDA graph per (batch size, probs) like the training
        step (four launches: conv1, conv2+pool, dense1, head)."""
        if not self.use_graph:
            self._forward(B, False)
            self._head(B, False, False, probs=probs)
            return
        key = ("inference", B, probs)
        g = self._graphs.get(key)
        if g is None:
            import gc
            gc.collect()
            torch.cuda.synchronize(self.dev)
            g = torch.cuda.CUDAGraph()
            cap = torch.cuda.Stream(device=self.dev)
            cap.wait_stream(torch.cuda.current_stream(self.dev))
            with torch.cuda.graph(g, stream=cap, capture_error_mode="thread_local"):
                self._forward(B, False)
                self._head(B, False, False, probs=probs)
            torch.cuda.current_stream(self.dev).wait_stream(cap)
            self._graphs[key] = g
        g.replay()

    def eval_step(self, x: torch.Tensor, y: torch.Tensor):
        B = x.shape[0]
        self.sync_params()
        if y.ndim == 2:
            y = y.argmax(dim=1)
        self.x_in[:B].copy_(x.reshape(B, -1), non_blocking=True)
        self.y_in[:B].copy_(y.to(torch.int64), non_blocking=True)
        self._inference(B, False)
        out = self.state_f[3].clone(), self.state_f[4].clone()
        self._ev_done[0].record(self._main())      # the next training batch staged into these buffers waits for us
        return out

    def predict(self, x: torch.Tensor):
        B = x.shape[0]
        self.sync_params()
        self.x_in[:B].copy_(x.reshape(B, -1), non_blocking=True)
        self.y_in[:B].zero_()
        self._inference(B, True)
        out = self.probs[:B].clone()
        self._ev_done[0].record(self._main())
        return out
