"""Keras-semantics optimizers operating on the flat arena.

Covers what the reference uses - ``keras.optimizers.Adadelta(1.0 * hvd.size())``
(``mnist_keras.py:84``) and ``tf.optimizers.Adam(0.001 * hvd.size())``
(``tensorflow2_keras_mnist.py:55``) - plus SGD (needed to exercise the LR-warmup
callback's momentum correction).

Math follows Keras, not torch.optim:

* Adadelta: rho=0.95, eps=1e-7.  ``a = rho*a + (1-rho)*g^2``;
  ``u = g*sqrt(d+eps)/sqrt(a+eps)``; ``p -= lr*u``; ``d = rho*d + (1-rho)*u^2``.
* Adam: b1=0.9, b2=0.999, eps=1e-7.  ``lr_t = lr*sqrt(1-b2^t)/(1-b1^t)``;
  ``m = b1*m+(1-b1)*g``; ``v = b2*v+(1-b2)*g^2``; ``p -= lr_t*m/(sqrt(v)+eps)``.

Because params/grads/slots are flat slabs, one update is a handful of whole-arena
elementwise ops here (the CPU oracle / generic path).  On B200 the same math is fused
into the peer-memory allreduce kernel (``csrc/comm/fused_allreduce.cu``), which reads
the hyper-parameters from the device-resident ``hyper`` vector kept by this class -
the learning rate is *data*, not a captured constant, so the warmup callback can change
it every batch without re-capturing the CUDA graph.
"""
from __future__ import annotations

import math
from typing import Dict, List

import torch

# ids shared with csrc/comm/fused_allreduce.cu
KIND_NONE, KIND_SGD, KIND_ADADELTA, KIND_ADAM = 0, 1, 2, 3

# layout of the device-side hyper-parameter vector (fp32[8])
H_LR, H_A, H_B, H_EPS, H_STEP, H_DECAY, H_LR_T, H_SPARE = range(8)


class Optimizer:
    kind = KIND_NONE
    slot_names: List[str] = []

    def __init__(self, lr: float, decay: float = 0.0, name: str = None, **kwargs):
        if "learning_rate" in kwargs:
            lr = kwargs.pop("learning_rate")
        if kwargs:
            raise TypeError(f"unexpected optimizer arguments: {sorted(kwargs)}")
        self._lr = float(lr)
        self.decay = float(decay)
        self.iterations = 0
        self.name = name or type(self).__name__
        self._hyper = None      # device fp32[2][8], created by bind()
        self._arena = None
        self._defer_push = False
        self._hyper_dirty = 0

    # -- learning rate is settable at run time (the warmup callback mutates it) --
    @property
    def lr(self) -> float:
        return self._lr

    @lr.setter
    def lr(self, v: float):
        self._lr = float(v)
        self._hyper_changed()

    def _hyper_changed(self) -> None:
        """A hyper-parameter was changed from the host.  Normally both device rows are rewritten now;
        the fused engine (which may still have the previous step's update kernel in flight) asks for
        the push to be deferred to the start of the next two steps, one row each."""
        if getattr(self, "_defer_push", False):
            self._hyper_dirty = 2
        else:
            self._push_hyper(both=True)

    learning_rate = lr

    @property
    def n_slots(self) -> int:
        return len(self.slot_names)

    def get_config(self) -> Dict:
        return {"lr": self._lr, "decay": self.decay}

    @classmethod
    def from_config(cls, config: Dict):
        return cls(**config)

    # -- binding to an arena ------------------------------------------------------
    def bind(self, arena) -> None:
        self._arena = arena
        # two rows: the fused engine alternates rows with its gradient-buffer parity, so the update
        # kernel of step t (possibly still running on a side stream) never sees step t+1's values
        self._hyper = torch.zeros(2, 8, dtype=torch.float32, device=arena.device)
        self._next_slot = 0
        # ring of pinned staging rows for the async H2D pushes; each entry carries the event of its last
        # copy, waited for before the entry is rewritten (the host may run many steps ahead of the device)
        self._pin_ring = [torch.zeros(8, dtype=torch.float32).pin_memory() for _ in range(16)] \
            if arena.device.type == "cuda" else None
        self._pin_ev = [None] * 16
        self._pin_np = None
        self._pin_i = 0
        self._push_hyper(both=True)

    def _effective_lr(self) -> float:
        lr = self._lr
        if self.decay > 0.0:
            lr = lr / (1.0 + self.decay * self.iterations)
        return lr

    def _lr_for_step(self) -> float:
        """LR for the step being applied (``iterations`` was already incremented)."""
        if self.decay > 0.0:
            return self._lr / (1.0 + self.decay * (self.iterations - 1))
        return self._lr

    def _hyper_values(self) -> List[float]:
        return [self._effective_lr(), 0.0, 0.0, 0.0, float(self.iterations), self.decay, 0.0, 0.0]

    def _push_hyper(self, both: bool = False) -> None:
        """Write the hyper-parameters the NEXT step will use into its row of the device vector."""
        if self._hyper is None:
            return
        vals = torch.tensor(self._hyper_values(), dtype=torch.float32)
        rows = (0, 1) if both else (self._next_slot,)
        if self._hyper.is_cuda:
            # tiny async H2D from a ring of pinned buffers; ordered on the current stream
            k = self._pin_i % len(self._pin_ring)
            buf = self._pin_ring[k]
            self._pin_i += 1
            ev = self._pin_ev[k]
            if ev is not None:
                ev.synchronize()          # the copy that last read this entry has executed (normally long ago)
            buf.copy_(vals)
            for r in rows:
                self._hyper[r].copy_(buf, non_blocking=True)
            if ev is None:
                ev = self._pin_ev[k] = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self._hyper.device))
        else:
            for r in rows:
                self._hyper[r].copy_(vals)

    def _push_hyper_raw(self, row: int, memcpy_async, stream) -> None:
        """Engine fast path: write this step's hyper-parameters into ``row`` with ONE 32-byte asynchronous copy on
        ``stream`` (the engine's input-staging copy stream, next to the batch's own H2D copies), so that a
        learning rate that changes every batch (warmup callback, Adam's lr_t) puts nothing on the compute stream.
        ``memcpy_async(dst, src, nbytes, raw_stream)`` is the C ABI's cudaMemcpyAsync wrapper."""
        k = self._pin_i % len(self._pin_ring)
        self._pin_i += 1
        ev = self._pin_ev[k]
        if ev is not None:
            ev.synchronize()
        buf = self._pin_ring[k]
        if self._pin_np is None:
            self._pin_np = [b.numpy() for b in self._pin_ring]
        self._pin_np[k][:] = self._hyper_values()
        rc = memcpy_async(self._hyper[row & 1].data_ptr(), buf.data_ptr(), 32, stream.cuda_stream)
        if rc:
            raise RuntimeError(f"hyper-parameter copy failed with CUDA error {rc}")
        if ev is None:
            ev = self._pin_ev[k] = torch.cuda.Event()
        ev.record(stream)

    @property
    def hyper(self) -> torch.Tensor:
        """Row 0 of the device hyper-parameter block (fp32[8]); row 1 is the engine's alternate."""
        return self._hyper[0]

    def hyper_row(self, which: int) -> torch.Tensor:
        return self._hyper[which & 1]

    # -- the update -----------------------------------------------------------------
    def apply_flat(self, params: torch.Tensor, grads: torch.Tensor, slots) -> None:
        raise NotImplementedError

    def step(self, arena=None, grads: torch.Tensor = None) -> None:
        """One update over the whole arena using already-averaged gradients."""
        arena = arena or self._arena
        g = arena.grads if grads is None else grads
        with torch.no_grad():
            self.iterations += 1
            self.apply_flat(arena.params, g, arena.slots)
        self._push_hyper()

    def get_gradients(self, arena) -> torch.Tensor:
        """Return the flat gradient slab to apply.  ``hvd.DistributedOptimizer`` overrides
        this to average across ranks first - the same hook Horovod overrides on Keras
        optimizers."""
        return arena.grads

    def apply_gradients(self, arena=None) -> None:
        arena = arena or self._arena
        self.step(arena, self.get_gradients(arena))

    # Keras state for checkpoints: list of (name, tensor) in Keras' weight order.
    def state_tensors(self, arena, layer_weight_names: List[str]):
        raise NotImplementedError

    def load_state_tensors(self, arena, layer_weight_names, tensors) -> None:
        raise NotImplementedError


class SGD(Optimizer):
    kind = KIND_SGD
    slot_names = ["momentum"]

    def __init__(self, lr=0.01, momentum=0.0, decay=0.0, nesterov=False, **kw):
        super().__init__(lr, decay, **kw)
        self.momentum = float(momentum)
        self.nesterov = bool(nesterov)

    def get_config(self):
        return {"lr": self._lr, "momentum": self.momentum, "decay": self.decay,
                "nesterov": self.nesterov}

    def _hyper_values(self):
        return [self._effective_lr(), self.momentum, 1.0 if self.nesterov else 0.0, 0.0,
                float(self.iterations), self.decay, 0.0, 0.0]

    def apply_flat(self, params, grads, slots):
        lr = self._lr_for_step()
        v = slots[0]
        v.mul_(self.momentum).add_(grads, alpha=-lr)            # v = mom*v - lr*g
        if self.nesterov:
            params.add_(v, alpha=self.momentum).add_(grads, alpha=-lr)
        else:
            params.add_(v)

    def state_tensors(self, arena, names):
        out = [("iterations", torch.tensor(self.iterations, dtype=torch.int64))]
        out += [(f"training/SGD/{n}/momentum", arena.slot(0, n)) for n in names]
        return out

    def load_state_tensors(self, arena, names, tensors):
        self.iterations = int(tensors[0].item())
        for n, t in zip(names, tensors[1:]):
            arena.slot(0, n).copy_(t.view_as(arena.slot(0, n)))


class Adadelta(Optimizer):
    kind = KIND_ADADELTA
    slot_names = ["accumulator", "delta_accumulator"]

    def __init__(self, lr=1.0, rho=0.95, epsilon=None, decay=0.0, **kw):
        super().__init__(lr, decay, **kw)
        self.rho = float(rho)
        self.epsilon = 1e-7 if epsilon is None else float(epsilon)

    def get_config(self):
        return {"lr": self._lr, "rho": self.rho, "decay": self.decay, "epsilon": self.epsilon}

    def _hyper_values(self):
        return [self._effective_lr(), self.rho, 0.0, self.epsilon, float(self.iterations),
                self.decay, 0.0, 0.0]

    def apply_flat(self, params, grads, slots):
        lr = self._lr_for_step()
        a, d = slots[0], slots[1]
        rho, eps = self.rho, self.epsilon
        a.mul_(rho).addcmul_(grads, grads, value=1.0 - rho)
        u = grads * torch.sqrt(d + eps) / torch.sqrt(a + eps)
        params.add_(u, alpha=-lr)
        d.mul_(rho).addcmul_(u, u, value=1.0 - rho)

    def state_tensors(self, arena, names):
        # Keras 2.x Adadelta.weights = accumulators + delta_accumulators (iterations is
        # not a saved weight for this optimizer in standalone Keras 2.2/2.3).
        out = [(f"training/Adadelta/{n}/accumulator", arena.slot(0, n)) for n in names]
        out += [(f"training/Adadelta/{n}/delta_accumulator", arena.slot(1, n)) for n in names]
        return out

    def load_state_tensors(self, arena, names, tensors):
        k = len(names)
        if len(tensors) == 2 * k + 1:      # a writer that also stored iterations first
            self.iterations = int(tensors[0].item())
            tensors = tensors[1:]
        for n, t in zip(names, tensors[:k]):
            arena.slot(0, n).copy_(t.view_as(arena.slot(0, n)))
        for n, t in zip(names, tensors[k:2 * k]):
            arena.slot(1, n).copy_(t.view_as(arena.slot(1, n)))


class Adam(Optimizer):
    kind = KIND_ADAM
    slot_names = ["m", "v"]

    def __init__(self, lr=0.001, beta_1=0.9, beta_2=0.999, epsilon=None, decay=0.0,
                 amsgrad=False, **kw):
        super().__init__(lr, decay, **kw)
        if amsgrad:
            raise NotImplementedError("amsgrad is not used by the reference")
        self.beta_1, self.beta_2 = float(beta_1), float(beta_2)
        self.epsilon = 1e-7 if epsilon is None else float(epsilon)

    def get_config(self):
        return {"lr": self._lr, "beta_1": self.beta_1, "beta_2": self.beta_2,
                "decay": self.decay, "epsilon": self.epsilon, "amsgrad": False}

    def lr_t(self, t: int) -> float:
        lr = self._lr
        if self.decay > 0:
            lr = lr / (1.0 + self.decay * (t - 1))
        return lr * math.sqrt(1.0 - self.beta_2 ** t) / (1.0 - self.beta_1 ** t)

    def _hyper_values(self):
        # H_LR_T is for the *next* step (t = iterations + 1); the fused kernel also
        # keeps its own device-side step counter so graph replays need no host push.
        t = self.iterations + 1
        return [self._effective_lr(), self.beta_1, self.beta_2, self.epsilon,
                float(self.iterations), self.decay, self.lr_t(t), 0.0]

    def apply_flat(self, params, grads, slots):
        t = self.iterations
        lr_t = self.lr_t(t)
        m, v = slots[0], slots[1]
        m.mul_(self.beta_1).add_(grads, alpha=1.0 - self.beta_1)
        v.mul_(self.beta_2).addcmul_(grads, grads, value=1.0 - self.beta_2)
        params.addcdiv_(m, torch.sqrt(v) + self.epsilon, value=-lr_t)

    def state_tensors(self, arena, names):
        out = [("training/Adam/iter", torch.tensor(self.iterations, dtype=torch.int64))]
        out += [(f"training/Adam/{n}/m", arena.slot(0, n)) for n in names]
        out += [(f"training/Adam/{n}/v", arena.slot(1, n)) for n in names]
        return out

    def load_state_tensors(self, arena, names, tensors):
        self.iterations = int(tensors[0].item())
        k = len(names)
        for n, t in zip(names, tensors[1:1 + k]):
            arena.slot(0, n).copy_(t.view_as(arena.slot(0, n)))
        for n, t in zip(names, tensors[1 + k:1 + 2 * k]):
            arena.slot(1, n).copy_(t.view_as(arena.slot(1, n)))


_REGISTRY = {"sgd": SGD, "adadelta": Adadelta, "adam": Adam}


def get(identifier):
    if isinstance(identifier, Optimizer):
        return identifier
    if isinstance(identifier, str):
        return _REGISTRY[identifier.lower()]()
    if isinstance(identifier, dict):
        return deserialize(identifier)
    raise ValueError(f"cannot interpret optimizer {identifier!r}")


def serialize(opt: Optimizer) -> Dict:
    return {"class_name": type(opt).__name__, "config": opt.get_config()}


def deserialize(cfg: Dict) -> Optimizer:
    cls = _REGISTRY[cfg["class_name"].lower()]
    c = dict(cfg.get("config", {}))
    if "learning_rate" in c:
        c["lr"] = c.pop("learning_rate")
    c.pop("name", None)
    return cls(**c)
