"""Losses and metrics with Keras semantics.

Reference use: ``keras.losses.categorical_crossentropy`` on one-hot targets
(``mnist_keras.py:89`` with ``to_categorical`` at ``:68-69``) and
``tf.losses.SparseCategoricalCrossentropy()`` on int labels over softmax
*probabilities* (``tensorflow2_keras_mnist.py:62``); metric ``'accuracy'``
(``mnist_keras.py:91``, ``tensorflow2_keras_mnist.py:64``).

Keras' TF backend works from probabilities: renormalise by the row sum, clip to
``[1e-7, 1-1e-7]``, then ``-sum(t*log p)`` averaged over the batch.
"""
from __future__ import annotations

import torch

EPSILON = 1e-7


def _as_probs(y_pred, from_logits):
    if from_logits:
        return torch.softmax(y_pred, dim=-1)
    p = y_pred / y_pred.sum(dim=-1, keepdim=True)
    return p


def categorical_crossentropy(y_true, y_pred, from_logits: bool = False):
    """Per-sample loss; ``y_true`` is one-hot (or soft) with the same shape as ``y_pred``."""
    p = _as_probs(y_pred, from_logits).clamp(EPSILON, 1.0 - EPSILON)
    return -(y_true.to(p.dtype) * torch.log(p)).sum(dim=-1)


def sparse_categorical_crossentropy(y_true, y_pred, from_logits: bool = False):
    p = _as_probs(y_pred, from_logits).clamp(EPSILON, 1.0 - EPSILON)
    idx = y_true.reshape(-1).to(torch.int64)
    flat = p.reshape(-1, p.shape[-1])
    return -torch.log(flat.gather(1, idx[:, None]).squeeze(1)).reshape(y_true.shape[: p.ndim - 1])


def mean_squared_error(y_true, y_pred):
    return ((y_pred - y_true.to(y_pred.dtype)) ** 2).mean(dim=-1)


class Loss:
    name = "loss"

    def __call__(self, y_true, y_pred):
        raise NotImplementedError


class CategoricalCrossentropy(Loss):
    name = "categorical_crossentropy"

    def __init__(self, from_logits=False):
        self.from_logits = from_logits

    def __call__(self, y_true, y_pred):
        return categorical_crossentropy(y_true, y_pred, self.from_logits)


class SparseCategoricalCrossentropy(Loss):
    name = "sparse_categorical_crossentropy"

    def __init__(self, from_logits=False):
        self.from_logits = from_logits

    def __call__(self, y_true, y_pred):
        return sparse_categorical_crossentropy(y_true, y_pred, self.from_logits)


class MeanSquaredError(Loss):
    name = "mean_squared_error"

    def __call__(self, y_true, y_pred):
        return mean_squared_error(y_true, y_pred)


_BY_NAME = {
    "categorical_crossentropy": categorical_crossentropy,
    "sparse_categorical_crossentropy": sparse_categorical_crossentropy,
    "mean_squared_error": mean_squared_error,
    "mse": mean_squared_error,
}


def get(identifier):
    if callable(identifier):
        return identifier
    return _BY_NAME[identifier]


def loss_name(fn) -> str:
    if isinstance(fn, Loss):
        return fn.name
    return getattr(fn, "__name__", str(fn))


def is_sparse(fn) -> bool:
    return "sparse" in loss_name(fn)


# ---------------------------------------------------------------- metrics
def categorical_accuracy(y_true, y_pred):
    return (y_true.argmax(dim=-1) == y_pred.argmax(dim=-1)).to(torch.float32)


def sparse_categorical_accuracy(y_true, y_pred):
    return (y_true.reshape(-1).to(torch.int64) == y_pred.argmax(dim=-1).reshape(-1)).to(torch.float32)


def resolve_accuracy(loss_fn, y_true, y_pred):
    """Keras resolves the string 'accuracy' from the loss / target shape."""
    if is_sparse(loss_fn) or y_true.ndim < y_pred.ndim or y_true.shape[-1] != y_pred.shape[-1]:
        return sparse_categorical_accuracy(y_true, y_pred)
    return categorical_accuracy(y_true, y_pred)
