"""Keras full-model HDF5 layout: ``model.save`` / ``load_model`` / ``load_weights``.

Reference use: ``ModelCheckpoint('checkpoint-{epoch}.h5')`` (``mnist_keras.py:101-104``,
``tensorflow2_keras_mnist.py:86-88``), ``model.save('keras-sample-model.h5')`` and
``load_model`` (``mnist_keras.py:118-119,124``).

Layout written:

* root attrs ``keras_version``, ``backend``, ``model_config`` (JSON), ``training_config`` (JSON)
* ``/model_weights`` attrs ``layer_names``, ``backend``, ``keras_version``; one sub-group per
  layer with attr ``weight_names``; datasets at ``/model_weights/<layer>/<layer>/kernel:0``
* ``/optimizer_weights`` attr ``weight_names`` + datasets (Adadelta accumulators / Adam
  iter, m, v) in Keras' ``optimizer.weights`` order.

Shapes are Keras': conv kernels HWIO, dense kernels (in, out), float32 little-endian.
Verified by our own reader and structural tests; not checked against real Keras/h5py
(neither exists on this image).
"""
from __future__ import annotations

import json
from typing import List

import numpy as np
import torch

from . import hdf5


def _np(t: torch.Tensor) -> np.ndarray:
    return t.detach().to("cpu").contiguous().numpy()


def _training_config(model) -> dict:
    from ..keras import losses as losses_mod, optimizers as opt_mod
    opt = model.optimizer
    return {
        "optimizer_config": opt_mod.serialize(opt),
        "loss": losses_mod.loss_name(model.loss) if model.loss is not None else None,
        "metrics": [m for m in model.metrics_names if m != "loss"],
        "sample_weight_mode": None,
        "loss_weights": None,
    }


def save_model(model, path: str, include_optimizer: bool = True, weights_only: bool = False) -> None:
    if not model.built:
        raise RuntimeError("model has no weights yet: call fit() or build() before save()")
    opt = model.optimizer
    if opt is not None and hasattr(opt, "materialize_state"):
        opt.materialize_state()          # gather sharded slots (two-shot owner-updates) to this rank
    with hdf5.Writer(path) as w:
        root_for_weights = "/" if weights_only else "/model_weights"
        if not weights_only:
            w.set_attr("/", "keras_version", model_keras_version())
            w.set_attr("/", "backend", b"tensorflow")
            cfg = {"class_name": type(model).__name__ if type(model).__name__ == "Sequential" else "Model",
                   "config": model.get_config()}
            w.set_attr("/", "model_config", json.dumps(cfg).encode("utf8"))
            if opt is not None and include_optimizer:
                w.set_attr("/", "training_config", json.dumps(_training_config(model)).encode("utf8"))
            w.create_group("/model_weights")
        lw = model._layer_weight_names()
        # one group per distinct layer name, in model order
        seen, layer_names = set(), []
        for lname, _ in lw:
            if lname not in seen:
                seen.add(lname)
                layer_names.append(lname)
        w.set_attr(root_for_weights, "layer_names", [n.encode() for n in layer_names])
        w.set_attr(root_for_weights, "backend", b"tensorflow")
        w.set_attr(root_for_weights, "keras_version", model_keras_version())
        per_layer = {n: [] for n in layer_names}
        for lname, wnames in lw:
            per_layer[lname] += wnames
        for lname in layer_names:
            g = f"{root_for_weights.rstrip('/')}/{lname}"
            w.create_group(g)
            wn = [f"{lname}/{s}:0" for s in per_layer[lname]]
            w.set_attr(g, "weight_names", [n.encode() for n in wn])
            for s in per_layer[lname]:
                full = f"{lname}/{s}" if f"{lname}/{s}" in model.arena._by_name else f"{lname}.{s}"
                w.create_dataset(f"{g}/{lname}/{s}:0", _np(model.arena.param(full)))
        if not weights_only and opt is not None and include_optimizer and opt.n_slots >= 0:
            names = model.trainable_weight_names()
            state = opt.state_tensors(model.arena, names)
            w.create_group("/optimizer_weights")
            w.set_attr("/optimizer_weights", "weight_names", [f"{n}:0".encode() for n, _ in state])
            for n, t in state:
                w.create_dataset(f"/optimizer_weights/{n}:0", _np(t))


def model_keras_version() -> bytes:
    from ..keras.models import KERAS_VERSION
    return KERAS_VERSION.encode()


def _weights_root(r: hdf5.Reader) -> str:
    return "/model_weights" if r.kind("/model_weights") == "group" else "/"


def load_weights_into(model, path: str) -> None:
    with hdf5.Reader(path) as r:
        root = _weights_root(r)
        layer_names = [b.decode() for b in r.attr(root, "layer_names")]
        have = dict(model._layer_weight_names_merged())
        with torch.no_grad():
            for lname in layer_names:
                g = f"{root.rstrip('/')}/{lname}"
                wn = [b.decode() for b in r.attr(g, "weight_names")]
                if not wn:
                    continue
                if lname not in have:
                    raise ValueError(f"checkpoint layer {lname!r} not present in the model")
                for full in wn:
                    short = full.split("/", 1)[1].rsplit(":", 1)[0]
                    arr = r.dataset(f"{g}/{full}")
                    key = f"{lname}/{short}" if f"{lname}/{short}" in model.arena._by_name else f"{lname}.{short}"
                    p = model.arena.param(key)
                    if tuple(arr.shape) != tuple(p.shape):
                        raise ValueError(f"shape mismatch for {full}: file {arr.shape} vs model {tuple(p.shape)}")
                    p.copy_(torch.from_numpy(arr))
    be = getattr(model, "_backend", None)
    if be is not None and hasattr(be, "weights_changed"):
        be.weights_changed()


def load_model(path: str, compile: bool = True, wrap_optimizer=None):
    """``wrap_optimizer``: optional callable applied to the deserialized optimizer before ``compile`` (``hvd.load_model``
    passes ``DistributedOptimizer``); the saved optimizer state is loaded into the wrapped instance."""
    from ..keras import models as models_mod, optimizers as opt_mod
    with hdf5.Reader(path) as r:
        attrs = r.attr_names("/")
        if "model_config" not in attrs:
            raise ValueError("no model_config in file (weights-only checkpoint?)")
        cfg = json.loads(r.attr("/", "model_config").decode("utf8"))
        if cfg["class_name"] != "Sequential":
            raise ValueError("load_model supports Sequential models")
        model = models_mod.Sequential.from_config(cfg["config"])
        tc = json.loads(r.attr("/", "training_config").decode("utf8")) if "training_config" in attrs else None
        first = cfg["config"]["layers"][0]["config"]
        model.build(tuple(first["batch_input_shape"][1:]))
        has_opt = r.kind("/optimizer_weights") == "group"
        opt_arrays: List[np.ndarray] = []
        if has_opt:
            for n in r.attr("/optimizer_weights", "weight_names"):
                opt_arrays.append(r.dataset("/optimizer_weights/" + n.decode()))
    load_weights_into(model, path)
    if compile and tc is not None:
        opt = opt_mod.deserialize(tc["optimizer_config"])
        if wrap_optimizer is not None:
            opt = wrap_optimizer(opt)
        model.compile(optimizer=opt, loss=tc["loss"], metrics=tc.get("metrics") or [])
        if has_opt and opt_arrays:
            with torch.no_grad():
                opt.load_state_tensors(model.arena, model.trainable_weight_names(),
                                       [torch.from_numpy(np.asarray(a)) for a in opt_arrays])
            opt._push_hyper()
    return model
