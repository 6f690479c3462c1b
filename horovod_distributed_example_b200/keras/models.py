"""Keras-like ``Sequential`` / ``TorchModel`` with ``compile / fit / evaluate / predict /
save / load_model``.

Covers the trainer surface the reference scripts use (``mnist_keras.py:71-119``,
``tensorflow2_keras_mnist.py:43-96``): ``Sequential`` (list or ``.add``), ``compile(loss,
optimizer, metrics)``, ``fit`` on arrays with ``batch_size``/``validation_data`` or on a
dataset with ``steps_per_epoch``, ``evaluate``, ``save`` and ``load_model``.

Two execution backends share this front end:

* ``_TorchBackend`` - plain torch autograd over the layer stack.  CPU oracle, and the
  generic GPU path for arbitrary models (ResNet-50 stress).  Gradients accumulate
  directly into the arena's gradient slab (``p.grad`` is an arena view).
* ``models.mnist_cnn.MnistEngine`` - the fused tcgen05/TMA CUDA-graph engine, selected
  automatically when the stack is the reference's MNIST CNN on a B200.

Both hand the arena's gradient slab to ``optimizer.apply_gradients``; with
``hvd.DistributedOptimizer`` that is where the cross-rank reduction happens.
"""
from __future__ import annotations

import collections
import json
import os
import time
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import callbacks as cb_mod
from . import layers as L
from . import losses as losses_mod
from . import optimizers as opt_mod
from .data import ArrayBatcher, Dataset
from ..parallel.arena import ParamArena

KERAS_VERSION = "2.2.4"
_name_counters: Dict[str, int] = {}


def clear_session():
    """Reset the automatic layer-name counters (``K.clear_session()``)."""
    _name_counters.clear()


def _auto_name(base: str) -> str:
    _name_counters[base] = _name_counters.get(base, 0) + 1
    return f"{base}_{_name_counters[base]}"


def _default_device() -> torch.device:
    from .. import basics
    if basics.is_initialized():
        return basics.device()
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() \
        else torch.device("cpu")


def _to_numpy(a):
    if isinstance(a, torch.Tensor):
        return a.detach().cpu().numpy()
    return np.asarray(a)


class _HostStager:
    """Moves one host batch to the device.  On CUDA it stages through two pinned buffers
    so the H2D copy is asynchronous and overlaps the previous step."""

    def __init__(self, device):
        self.device = device
        self._pinned = {}
        self._flip = 0

    def __call__(self, arr: np.ndarray, dtype=None) -> torch.Tensor:
        t = torch.from_numpy(np.ascontiguousarray(arr))
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        if self.device.type != "cuda":
            return t
        key = (tuple(t.shape), t.dtype, self._flip)
        buf = self._pinned.get(key)
        if buf is None:
            buf = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            self._pinned[key] = buf
        buf.copy_(t)
        return buf.to(self.device, non_blocking=True)

    def next_step(self):
        self._flip ^= 1


class _Resolved:
    def __init__(self, value):
        self._v = value

    def result(self):
        return self._v


class Model:
    """Trainer front end.  Subclasses provide ``_weight_specs`` and ``_forward``."""

    def __init__(self, name: Optional[str] = None):
        self.name = name or _auto_name("sequential")
        self.optimizer: Optional[opt_mod.Optimizer] = None
        self.loss = None
        self.metrics_names: List[str] = ["loss"]
        self.stop_training = False
        self.history = None
        self.arena: Optional[ParamArena] = None
        self.device: Optional[torch.device] = None
        self._backend = None
        self._compiled = False
        self._seed = None
        self.built = False

    # ---- to be provided by subclasses ---------------------------------------------
    def _build(self, input_shape): raise NotImplementedError
    def _forward(self, x, training, rng=None): raise NotImplementedError
    def _layer_weight_names(self) -> List[Tuple[str, List[str]]]: raise NotImplementedError
    def get_config(self) -> Dict: raise NotImplementedError

    # ---- arena ----------------------------------------------------------------------
    def _make_arena(self, specs, inits, device, n_slots=0):
        # Always a plain local arena here: building a model is a rank-local act (the reference loads a
        # model on rank 0 only, mnist_keras.py:124).  compile() with a DistributedOptimizer - which
        # every rank executes - is what moves it into peer-mapped symmetric memory (a collective).
        self.arena = ParamArena(specs, device, n_slots)
        g = torch.Generator().manual_seed(self._seed) if self._seed is not None else None
        with torch.no_grad():
            for (name, _shape), init in zip(specs, inits):
                self.arena.param(name).copy_(init(g).to(torch.float32))
        self.device = device

    def _ensure_arena(self, n_slots: int, symmetric: bool):
        """Re-home the arena when the optimizer needs more slots or symmetric (peer-mapped) memory."""
        ar = self.arena
        if ar.n_slots >= n_slots and (ar.symmetric or not symmetric):
            return
        allocator, double = None, ar.n_grad_bufs == 2
        if symmetric or ar.symmetric:
            from .. import basics
            from ..parallel import peer
            allocator = peer.context().alloc_tensor
            double = True                    # double-buffered grads: no one-shot exit barrier, and the big
                                             # bucket's update may overlap the next step's forward
        self.arena = ar.resize_slots(max(n_slots, ar.n_slots), allocator, double_grads=double)
        self._rebind_weights()

    def _rebind_weights(self): raise NotImplementedError

    # ---- compile ----------------------------------------------------------------
    def compile(self, optimizer="sgd", loss=None, metrics=None, **kwargs):
        """``experimental_run_tf_function`` and other TF-only kwargs are accepted and
        ignored (``tensorflow2_keras_mnist.py:62-65``): here the optimizer wrapper is
        always on the gradient path, so there is nothing to opt into."""
        self.optimizer = opt_mod.get(optimizer)
        self.loss = losses_mod.get(loss) if loss is not None else None
        self._metric_accuracy = False
        self.metrics_names = ["loss"]
        for m in metrics or []:
            if m in ("accuracy", "acc"):
                self._metric_accuracy = True
                self.metrics_names.append("accuracy")
            else:
                raise ValueError(f"unsupported metric {m!r} (the reference uses 'accuracy')")
        self._compiled = True
        self._backend = None
        if self.built:
            self._finish_compile()

    def _finish_compile(self):
        from .. import basics
        want_symm = False
        if self.device.type == "cuda" and basics.is_initialized() and getattr(self.optimizer, "_hvd_distributed", False):
            from ..parallel import peer
            want_symm = peer.available()
        self._ensure_arena(self.optimizer.n_slots, want_symm)
        self.optimizer.bind(self.arena)
        if hasattr(self.optimizer, "attach_model"):
            self.optimizer.attach_model(self)
        self._backend = self._select_backend()

    def _select_backend(self):
        if self.device.type == "cuda":
            from ..models import mnist_cnn
            eng = mnist_cnn.maybe_engine(self)
            if eng is not None:
                return eng
            # Not silent: everything below runs on stock torch ops (cuDNN/cuBLAS); only the optimizer path
            # (arena + fused allreduce/update buckets) is this framework's own code for such models.
            from .. import basics
            basics.log.warning("model %s does not match the fused sm_100a MNIST engine (%s); its forward/backward "
                               "run on torch autograd (cuDNN/cuBLAS)", self.name, mnist_cnn.why_not(self))
        return _TorchBackend(self)

    def _ensure_built(self, x_shape):
        if not self.built:
            self._build(tuple(x_shape[1:]))
        if self._compiled and self._backend is None:
            self._finish_compile()

    # ---- weights ------------------------------------------------------------------
    def _sync_backend(self):
        """Wait for any parameter update the backend still has in flight on a side stream."""
        be = self._backend
        if be is not None and hasattr(be, "sync_params"):
            be.sync_params()

    def get_weights(self) -> List[np.ndarray]:
        self._sync_backend()
        return [self.arena.param(n).detach().cpu().numpy().copy() for n in self.trainable_weight_names()]

    def set_weights(self, weights: Sequence[np.ndarray]):
        self._sync_backend()
        it = iter(weights)
        with torch.no_grad():
            for n in self.trainable_weight_names():
                p = self.arena.param(n)
                p.copy_(torch.as_tensor(next(it)).view_as(p))
        if self._backend is not None and hasattr(self._backend, "weights_changed"):
            self._backend.weights_changed()

    def count_params(self) -> int:
        return self.arena.n_params()

    def _layer_weight_names_merged(self):
        merged = {}
        for lname, ws in self._layer_weight_names():
            merged.setdefault(lname, []).extend(ws)
        return list(merged.items())

    def trainable_weight_names(self) -> List[str]:
        return [f"{ln}/{w}" for ln, ws in self._layer_weight_names() for w in ws]

    # ---- steps ----------------------------------------------------------------------
    def _prep_y(self, y: np.ndarray):
        return y

    def train_on_batch(self, x, y):
        """One optimisation step on a host batch.  Returns ``[loss, accuracy]`` floats."""
        x, y = _to_numpy(x), _to_numpy(y)
        self._ensure_built(x.shape)
        if hasattr(self._backend, "train_host_batch"):
            loss, acc = self._backend.train_host_batch(x, y)     # direct H2D into the engine's buffers
            return [loss, acc] if self._metric_accuracy else loss
        st = self._stager()
        xd, yd = st(x, torch.float32), st(y)
        st.next_step()
        loss, acc = self._backend.train_step(xd, yd)
        out = [float(loss)]
        if self._metric_accuracy:
            out.append(float(acc))
        return out if len(out) > 1 else out[0]

    def train_on_batch_async(self, x, y):
        """Non-blocking ``train_on_batch``: enqueues the host->device copy of this batch, the step and the
        device->host read of its (loss, accuracy), and returns a handle whose ``result()`` waits for that
        step only.  Resolving handles one step late lets the copy of batch t+1 overlap the compute of
        step t (the fused engine double-buffers its inputs); other backends resolve immediately."""
        x, y = _to_numpy(x), _to_numpy(y)
        self._ensure_built(x.shape)
        if hasattr(self._backend, "train_host_batch_async"):
            return self._backend.train_host_batch_async(x, y)
        out = self.train_on_batch(x, y)
        return _Resolved(tuple(out) if isinstance(out, list) else (out, None))

    def test_on_batch(self, x, y):
        x, y = _to_numpy(x), _to_numpy(y)
        self._ensure_built(x.shape)
        st = self._stager()
        ls, cs = self._backend.eval_step(st(x, torch.float32), st(y))
        n = x.shape[0]
        out = [float(ls) / n]
        if self._metric_accuracy:
            out.append(float(cs) / n)
        return out if len(out) > 1 else out[0]

    def _stager(self):
        if getattr(self, "_stage", None) is None or self._stage.device != self.device:
            self._stage = _HostStager(self.device)
        return self._stage

    # ---- fit ------------------------------------------------------------------------
    def fit(self, x=None, y=None, batch_size=None, epochs=1, verbose=1, callbacks=None,
            validation_data=None, shuffle=True, initial_epoch=0, steps_per_epoch=None,
            validation_steps=None, **kwargs):
        if not self._compiled:
            raise RuntimeError("You must compile your model before training.")
        if isinstance(x, Dataset):
            it_factory, steps, samples, bs = self._dataset_epochs(x, steps_per_epoch)
            first = next(iter(x))
            self._ensure_built(np.asarray(first[0]).shape)
        else:
            x, y = _to_numpy(x), _to_numpy(y)
            bs = batch_size or 32
            batcher = ArrayBatcher([x, y], bs, shuffle, seed=self._seed)
            steps = steps_per_epoch or batcher.steps
            samples = x.shape[0]
            self._ensure_built(x.shape)

            def it_factory():
                return batcher.epoch()

        do_val = validation_data is not None
        metrics = list(self.metrics_names) + (["val_" + m for m in self.metrics_names] if do_val else [])
        params = {"epochs": epochs, "steps": steps, "samples": samples, "batch_size": bs,
                  "verbose": verbose, "do_validation": do_val, "metrics": metrics}
        self.history = cb_mod.History()
        cbs = list(callbacks or []) + [self.history]
        if verbose:
            cbs.append(cb_mod.ProgbarLogger())
        cl = cb_mod.CallbackList(cbs, self, params)
        self.stop_training = False
        st = self._stager()
        be = self._backend

        cl.on_train_begin({})
        stream = None
        for epoch in range(initial_epoch, epochs):
            cl.on_epoch_begin(epoch, {})
            # No host sync per step.  A backend that publishes step results asynchronously
            # (the fused engine: a pinned host ring) gives handles that are summed on the host as they land, a
            # few steps behind the launches; other backends return device scalars summed on the device.
            asynchronous = hasattr(be, "train_host_batch_async")
            loss_sum = torch.zeros((), dtype=torch.float64, device=self.device)
            acc_sum = torch.zeros((), dtype=torch.float64, device=self.device)
            loss_h = acc_h = 0.0
            pend = collections.deque()
            seen = 0
            if stream is None or not isinstance(x, Dataset):
                stream = iter(it_factory())
            for step in range(steps):
                try:
                    xb, yb = next(stream)
                except StopIteration:
                    break
                n = len(xb)
                cl.on_batch_begin(step, {"batch": step, "size": n})
                if asynchronous:
                    h = be.train_host_batch_async(xb, yb)        # engine stages the batch itself (pipelined H2D)
                    pend.append((h, n))
                    loss, acc = h.lazy()
                    while pend and (len(pend) > 8 or pend[0][0].ready()):
                        hh, nn = pend.popleft()
                        lv, av = hh.result()
                        loss_h += lv * nn
                        acc_h += av * nn
                else:
                    xd, yd = st(xb, torch.float32), st(yb)
                    st.next_step()
                    loss, acc = be.train_step(xd, yd)
                    loss_sum += loss.detach().to(torch.float64) * n
                    if self._metric_accuracy:
                        acc_sum += acc.detach().to(torch.float64) * n
                seen += n
                logs = {"batch": step, "size": n, "loss": loss}
                if self._metric_accuracy:
                    logs["accuracy"] = acc
                cl.on_batch_end(step, logs)
                if self.stop_training:
                    break
            while pend:
                hh, nn = pend.popleft()
                lv, av = hh.result()
                loss_h += lv * nn
                acc_h += av * nn
            if hasattr(be, "flush_params"):
                be.flush_params()        # every rank: rows it owns reach the other replicas even if training stops here
            elogs = {"loss": (loss_h if asynchronous else float(loss_sum)) / max(seen, 1)}
            if self._metric_accuracy:
                elogs["accuracy"] = (acc_h if asynchronous else float(acc_sum)) / max(seen, 1)
            if do_val:
                vx, vy = validation_data[0], validation_data[1]
                v = self.evaluate(vx, vy, batch_size=bs, verbose=0)
                v = v if isinstance(v, list) else [v]
                for name, val in zip(self.metrics_names, v):
                    elogs["val_" + name] = val
            cl.on_epoch_end(epoch, elogs)
            if self.stop_training:
                break
        cl.on_train_end({})
        return self.history

    def _dataset_epochs(self, ds: Dataset, steps_per_epoch):
        card = ds.cardinality
        if steps_per_epoch is None:
            if card < 0:
                raise ValueError("steps_per_epoch is required for an infinite dataset")
            steps_per_epoch = card
        bs = None
        for op, arg in ds._ops:
            if op == "batch":
                bs = arg[0]

        def factory():
            return iter(ds)
        return factory, steps_per_epoch, None, bs

    # ---- evaluate / predict ------------------------------------------------------
    def evaluate(self, x=None, y=None, batch_size=None, verbose=1, steps=None, **kwargs):
        if isinstance(x, Dataset):
            batches = iter(x)
            first_shape = None
        else:
            x, y = _to_numpy(x), _to_numpy(y)
            bs = batch_size or 32
            batches = ArrayBatcher([x, y], bs, shuffle=False).epoch()
            first_shape = x.shape
        if first_shape is not None:
            self._ensure_built(first_shape)
        st = self._stager()
        # per-batch (loss sum, correct count) stay on the device and are reduced once at the end: no host sync and
        # no extra kernels per batch
        parts_l, parts_c = [], []
        seen = 0
        for i, (xb, yb) in enumerate(batches):
            if steps is not None and i >= steps:
                break
            if first_shape is None:
                self._ensure_built(np.asarray(xb).shape)
                first_shape = xb.shape
            xd, yd = st(xb, torch.float32), st(yb)
            st.next_step()
            ls, cs = self._backend.eval_step(xd, yd)
            parts_l.append(ls.detach().reshape(()))
            parts_c.append(cs.detach().reshape(()))
            seen += len(xb)
        if not parts_l:
            out = [0.0] + ([0.0] if self._metric_accuracy else [])
            return out if len(out) > 1 else out[0]
        loss_sum = torch.stack(parts_l).to(torch.float64).sum()
        out = [float(loss_sum) / max(seen, 1)]
        if self._metric_accuracy:
            out.append(float(torch.stack(parts_c).to(torch.float64).sum()) / max(seen, 1))
        return out if len(out) > 1 else out[0]

    def predict(self, x, batch_size=32, **kwargs):
        x = _to_numpy(x)
        self._ensure_built(x.shape)
        if self._backend is None:
            self._backend = _TorchBackend(self)
        st = self._stager()
        outs = []
        for (xb,) in ArrayBatcher([x], batch_size, shuffle=False).epoch():
            outs.append(self._backend.predict(st(xb, torch.float32)).detach().cpu())
            st.next_step()
        return torch.cat(outs).numpy()

    def __call__(self, x, training=False):
        return self._forward(x, training)

    # ---- persistence ----------------------------------------------------------------
    def save(self, filepath, overwrite=True, include_optimizer=True):
        from ..utils import keras_h5
        if os.path.exists(filepath) and not overwrite:
            raise FileExistsError(filepath)
        self._sync_backend()
        keras_h5.save_model(self, filepath, include_optimizer=include_optimizer)

    def save_weights(self, filepath, overwrite=True):
        from ..utils import keras_h5
        self._sync_backend()
        keras_h5.save_model(self, filepath, include_optimizer=False, weights_only=True)

    def load_weights(self, filepath):
        from ..utils import keras_h5
        self._sync_backend()
        keras_h5.load_weights_into(self, filepath)

    def summary(self, print_fn=print):
        print_fn(f'Model: "{self.name}"')
        print_fn("_" * 65)
        print_fn(f"{'Layer (type)':<29}{'Output Shape':<26}{'Param #':<10}")
        print_fn("=" * 65)
        for lay in getattr(self, "layers", []):
            shp = str((None,) + tuple(lay.output_shape)) if lay.output_shape else "?"
            print_fn(f"{lay.name + ' (' + type(lay).__name__ + ')':<29}{shp:<26}{lay.count_params():<10}")
        print_fn("=" * 65)
        print_fn(f"Total params: {self.count_params():,}")


class Sequential(Model):
    def __init__(self, layers: Optional[Sequence[L.Layer]] = None, name: Optional[str] = None,
                 seed: Optional[int] = None):
        super().__init__(name)
        self.layers: List[L.Layer] = []
        self._seed = seed
        for lay in layers or []:
            self.add(lay)

    def add(self, layer: L.Layer):
        if self.built:
            raise RuntimeError("cannot add layers after the model is built")
        if layer.name is None:
            layer.name = _auto_name(layer.base_name)
        self.layers.append(layer)
        # Keras builds eagerly when the first layer carries input_shape (mnist_keras.py:72-74)
        return self

    def build(self, input_shape=None):
        if input_shape is None:
            input_shape = self.layers[0].input_shape_arg
        if input_shape is None:
            raise ValueError("input shape unknown: pass input_shape to the first layer or call fit")
        if len(input_shape) and input_shape[0] is None:
            input_shape = input_shape[1:]
        self._build(tuple(input_shape))

    def _build(self, input_shape):
        declared = self.layers[0].input_shape_arg
        if declared is not None and tuple(declared) != tuple(input_shape):
            raise ValueError(f"input shape {input_shape} does not match declared {declared}")
        device = _default_device()
        per_layer = []
        shape = tuple(input_shape)
        for lay in self.layers:
            specs = lay.build(shape)
            per_layer.append(specs)
            shape = lay.output_shape
        # arena order = backward-ready order: last layer first; bias before kernel
        # (SURVEY.md K-C1 message order)
        specs, inits = [], []
        for lay, lspecs in reversed(list(zip(self.layers, per_layer))):
            for short, shp, init in reversed(lspecs):
                specs.append((f"{lay.name}/{short}", shp))
                inits.append(init)
        self._per_layer_specs = per_layer
        n_slots = self.optimizer.n_slots if self.optimizer is not None else 0
        self._make_arena(specs, inits, device, n_slots)
        self.input_shape = tuple(input_shape)
        self.output_shape = shape
        self._rebind_weights()
        self.built = True

    def _rebind_weights(self):
        for lay, lspecs in zip(self.layers, self._per_layer_specs):
            lay.weights = {}
            for short, _shp, _init in lspecs:
                name = f"{lay.name}/{short}"
                w = self.arena.param(name)
                w.requires_grad_(True)
                w.grad = self.arena.grad(name)
                lay.weights[short] = w

    def _forward(self, x, training, rng=None):
        for lay in self.layers:
            x = lay.call(x, training, rng)
        return x

    def _layer_weight_names(self):
        return [(lay.name, lay.weight_names()) for lay in self.layers]

    def get_config(self):
        cfg_layers = []
        for i, lay in enumerate(self.layers):
            c = lay.get_config()
            if i == 0 and self.built:
                c["batch_input_shape"] = [None] + list(self.input_shape)
                c["dtype"] = "float32"
            cfg_layers.append({"class_name": type(lay).__name__, "config": c})
        return {"name": self.name, "layers": cfg_layers}

    @classmethod
    def from_config(cls, config, seed=None):
        m = cls(name=config.get("name"), seed=seed)
        for lc in config["layers"]:
            m.add(L.from_config(lc["class_name"], lc["config"]))
        return m


class TorchModel(Model):
    """Wrap any ``torch.nn.Module`` so it trains through the same arena /
    ``DistributedOptimizer`` path (used for the ResNet-50 large-bucket stress,
    BASELINE.json config 5).  The module's parameters are re-homed into the arena in
    reverse registration order (approximately backward-ready order, as DDP assumes)."""

    def __init__(self, module: torch.nn.Module, name: Optional[str] = None, channels_last_input=True):
        super().__init__(name or _auto_name("torch_model"))
        self.module = module
        self.channels_last_input = channels_last_input
        self.layers = []

    def _build(self, input_shape):
        device = _default_device()
        self.module.to(device)
        named = [(n, p) for n, p in self.module.named_parameters() if p.requires_grad]
        specs = [(n, tuple(p.shape)) for n, p in reversed(named)]
        inits = [(lambda g, p=p: p.detach().cpu()) for _n, p in reversed(named)]
        n_slots = self.optimizer.n_slots if self.optimizer is not None else 0
        self._make_arena(specs, inits, device, n_slots)
        self._named = named
        self._rebind_weights()
        self.input_shape = tuple(input_shape)
        self.built = True

    def _rebind_weights(self):
        for n, p in self._named:
            p.data = self.arena.param(n)
            p.grad = self.arena.grad(n)

    def _forward(self, x, training, rng=None):
        self.module.train(training)
        if self.channels_last_input and x.ndim == 4:
            x = x.permute(0, 3, 1, 2)
        return self.module(x)

    def _layer_weight_names(self):
        return [(n.rsplit(".", 1)[0] if "." in n else n, [n.rsplit(".", 1)[-1]]) for n, _ in self._named]

    def trainable_weight_names(self):
        return [n for n, _ in self._named]

    def get_config(self):
        return {"name": self.name, "layers": [], "torch_module": type(self.module).__name__}


class _TorchBackend:
    """Autograd execution of the layer stack; grads land in the arena slab."""

    name = "torch-autograd"

    def __init__(self, model: Model):
        self.m = model
        self.rng = None
        if model._seed is not None:
            self.rng = torch.Generator(device=model.device).manual_seed(model._seed + 1)

    def _loss_acc(self, y, y_pred):
        per = self.m.loss(y, y_pred)
        acc = losses_mod.resolve_accuracy(self.m.loss, y, y_pred) if self.m._metric_accuracy else None
        return per, acc

    def train_step(self, x, y):
        m = self.m
        arena = m.arena
        opt = m.optimizer
        if getattr(opt, "_hooked", False):
            # a fused bucket kernel of the previous step may still be reading the gradient slab on the side stream
            torch.cuda.current_stream(m.device).wait_stream(opt._fused.side_stream)
        # backward_passes_per_step > 1: gradients keep accumulating in the slab over the passes of a cycle
        # (the wrapper only reduces + applies on the last one), so the slab is cleared on the first pass only
        if getattr(opt, "_pass", 0) % getattr(opt, "_bpps", 1) == 0:
            arena.grads.zero_()
        if hasattr(opt, "begin_step"):
            opt.begin_step()
        with torch.enable_grad():
            y_pred = m._forward(x, True, self.rng)
            per, acc = self._loss_acc(y, y_pred)
            loss = per.mean()
        loss.backward()
        opt.apply_gradients(arena)
        return loss.detach(), (acc.mean() if acc is not None else None)

    def eval_step(self, x, y):
        with torch.no_grad():
            y_pred = self.m._forward(x, False)
            per, acc = self._loss_acc(y, y_pred)
            return per.sum(), (acc.sum() if acc is not None else torch.zeros((), device=x.device))

    def predict(self, x):
        with torch.no_grad():
            return self.m._forward(x, False)


def load_model(filepath, custom_objects=None, compile=True):
    """Load a full-model ``.h5`` written by ``Model.save`` (reference: ``mnist_keras.py:124``)."""
    from ..utils import keras_h5
    return keras_h5.load_model(filepath, compile=compile)
