"""Keras-like front end (``Sequential``, layers, optimizers, losses, callbacks, datasets)
- the L2 layer of the reference stack, rebuilt over the arena."""
from . import backend, callbacks, datasets, layers, losses, optimizers, utils  # noqa: F401
from .data import Dataset  # noqa: F401
from .models import Model, Sequential, TorchModel, load_model  # noqa: F401
