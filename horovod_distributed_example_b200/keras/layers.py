"""Keras-like layers (NHWC, Keras weight layouts) for the generic / oracle path.

The reference model (``mnist_keras.py:71-81``, ``tensorflow2_keras_mnist.py:43-52``) is
``Conv2D(32,3x3,relu) -> Conv2D(64,3x3,relu) -> MaxPooling2D(2x2) -> Dropout(0.25) ->
Flatten -> Dense(128,relu) -> Dropout(0.5) -> Dense(10,softmax)``.

Weights are kept in Keras' on-disk layouts - conv kernels HWIO, dense kernels
``(in, out)``, Flatten order H,W,C - so a checkpoint is a plain dump of the arena and
``dense_1/kernel`` row order matches what Keras expects.

``call()`` here is plain torch (autograd) and serves as the CPU oracle and as the generic
GPU path for arbitrary stacks.  When the stack is exactly the MNIST CNN on a B200, the
model swaps in the fused tcgen05 engine (``models/mnist_cnn.py``) instead.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

_ACTIVATIONS = {
    None: lambda x: x,
    "linear": lambda x: x,
    "relu": F.relu,
    "softmax": lambda x: F.softmax(x, dim=-1),
    "sigmoid": torch.sigmoid,
    "tanh": torch.tanh,
}


def _pair(v):
    if isinstance(v, (tuple, list)):
        return int(v[0]), int(v[1])
    return int(v), int(v)


def glorot_uniform(shape, fan_in, fan_out, generator=None):
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    return (torch.rand(shape, generator=generator) * 2.0 - 1.0) * limit


# keras.backend.image_data_format(): 'channels_last' (default, what the fused engine runs) or 'channels_first'
# (the reference's other input branch, mnist_keras.py:50-53; runs on the torch-autograd backend)
_image_data_format = "channels_last"


def image_data_format() -> str:
    return _image_data_format


def set_image_data_format(fmt: str) -> None:
    global _image_data_format
    if fmt not in ("channels_last", "channels_first"):
        raise ValueError("data_format must be 'channels_last' or 'channels_first'")
    _image_data_format = fmt


def _fmt(data_format):
    f = data_format or _image_data_format
    if f not in ("channels_last", "channels_first"):
        raise ValueError("data_format must be 'channels_last' or 'channels_first'")
    return f


class Layer:
    """Base layer.  Subclasses define ``build`` (returns weight specs + output shape)
    and ``call``."""

    _counters: Dict[str, int] = {}
    base_name = "layer"

    def __init__(self, name: Optional[str] = None, input_shape=None, **kwargs):
        if kwargs:
            raise TypeError(f"{type(self).__name__}: unexpected arguments {sorted(kwargs)}")
        self._given_name = name
        self.name = name
        self.input_shape_arg = tuple(input_shape) if input_shape is not None else None
        self.built = False
        self.weights: Dict[str, torch.Tensor] = {}      # short name -> tensor view
        self.input_shape = None
        self.output_shape = None

    # weight specs: list of (short_name, shape, initializer_fn(generator)->tensor)
    def build(self, input_shape: Tuple[int, ...]):
        self.input_shape = tuple(input_shape)
        self.output_shape = self.compute_output_shape(input_shape)
        self.built = True
        return []

    def compute_output_shape(self, input_shape):
        return tuple(input_shape)

    def call(self, x, training: bool, rng=None):
        return x

    def weight_names(self) -> List[str]:
        return list(self.weights.keys())

    def get_config(self) -> Dict:
        return {"name": self.name, "trainable": True}

    def count_params(self) -> int:
        return sum(w.numel() for w in self.weights.values())


class InputLayer(Layer):
    base_name = "input"


class Conv2D(Layer):
    base_name = "conv2d"

    def __init__(self, filters, kernel_size, strides=(1, 1), padding="valid", activation=None,
                 use_bias=True, data_format=None, **kw):
        super().__init__(**kw)
        self.data_format = _fmt(data_format)
        self.filters = int(filters)
        self.kernel_size = _pair(kernel_size)
        self.strides = _pair(strides)
        self.padding = padding.lower()
        if self.padding not in ("valid", "same"):
            raise ValueError("padding must be 'valid' or 'same'")
        self.activation = activation
        self.use_bias = use_bias

    def compute_output_shape(self, s):
        cf = self.data_format == "channels_first"
        h, w = (s[1], s[2]) if cf else (s[0], s[1])
        kh, kw = self.kernel_size
        sh, sw = self.strides
        if self.padding == "valid":
            oh, ow = (h - kh) // sh + 1, (w - kw) // sw + 1
        else:
            oh, ow = math.ceil(h / sh), math.ceil(w / sw)
        return (self.filters, oh, ow) if cf else (oh, ow, self.filters)

    def build(self, input_shape):
        super().build(input_shape)
        cin = input_shape[0] if self.data_format == "channels_first" else input_shape[-1]
        kh, kw = self.kernel_size
        fan_in, fan_out = kh * kw * cin, kh * kw * self.filters
        specs = [("kernel", (kh, kw, cin, self.filters),
                  lambda g: glorot_uniform((kh, kw, cin, self.filters), fan_in, fan_out, g))]
        if self.use_bias:
            specs.append(("bias", (self.filters,), lambda g: torch.zeros(self.filters)))
        return specs

    def call(self, x, training, rng=None):
        # x: (N,H,W,C) -> NCHW view; kernel HWIO -> OIHW view (no copies on the hot path
        # matter here: this is the oracle path)
        w = self.weights["kernel"].permute(3, 2, 0, 1)
        b = self.weights.get("bias")
        cf = self.data_format == "channels_first"
        xin = x if cf else x.permute(0, 3, 1, 2)
        if self.padding == "same":
            kh, kw = self.kernel_size
            ph, pw = kh - 1, kw - 1
            xin = F.pad(xin, (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2))
        y = F.conv2d(xin, w, b, stride=self.strides)
        if not cf:
            y = y.permute(0, 2, 3, 1)
        return _ACTIVATIONS[self.activation](y)

    def get_config(self):
        return {"name": self.name, "trainable": True, "filters": self.filters,
                "kernel_size": list(self.kernel_size), "strides": list(self.strides),
                "padding": self.padding, "data_format": self.data_format,
                "dilation_rate": [1, 1], "activation": self.activation or "linear",
                "use_bias": self.use_bias,
                "kernel_initializer": {"class_name": "VarianceScaling", "config": {
                    "scale": 1.0, "mode": "fan_avg", "distribution": "uniform", "seed": None}},
                "bias_initializer": {"class_name": "Zeros", "config": {}},
                "kernel_regularizer": None, "bias_regularizer": None,
                "activity_regularizer": None, "kernel_constraint": None, "bias_constraint": None}


class MaxPooling2D(Layer):
    base_name = "max_pooling2d"

    def __init__(self, pool_size=(2, 2), strides=None, padding="valid", data_format=None, **kw):
        super().__init__(**kw)
        self.data_format = _fmt(data_format)
        self.pool_size = _pair(pool_size)
        self.strides = _pair(strides) if strides is not None else self.pool_size
        self.padding = padding.lower()

    def compute_output_shape(self, s):
        cf = self.data_format == "channels_first"
        c, h, w = (s[0], s[1], s[2]) if cf else (s[2], s[0], s[1])
        oh = (h - self.pool_size[0]) // self.strides[0] + 1
        ow = (w - self.pool_size[1]) // self.strides[1] + 1
        return (c, oh, ow) if cf else (oh, ow, c)

    _pool = staticmethod(F.max_pool2d)

    def call(self, x, training, rng=None):
        if self.data_format == "channels_first":
            return self._pool(x, self.pool_size, self.strides)
        y = self._pool(x.permute(0, 3, 1, 2), self.pool_size, self.strides)
        return y.permute(0, 2, 3, 1)

    def get_config(self):
        return {"name": self.name, "trainable": True, "pool_size": list(self.pool_size),
                "padding": self.padding, "strides": list(self.strides),
                "data_format": self.data_format}


class AveragePooling2D(MaxPooling2D):
    base_name = "average_pooling2d"
    _pool = staticmethod(F.avg_pool2d)


class GlobalAveragePooling2D(Layer):
    base_name = "global_average_pooling2d"

    def compute_output_shape(self, s):
        return (s[-1],)

    def call(self, x, training, rng=None):
        return x.mean(dim=(1, 2))


class Dropout(Layer):
    base_name = "dropout"

    def __init__(self, rate, **kw):
        super().__init__(**kw)
        self.rate = float(rate)
        if not 0.0 <= self.rate < 1.0:
            raise ValueError("dropout rate must be in [0, 1)")

    def call(self, x, training, rng=None):
        if not training or self.rate == 0.0:
            return x
        keep = 1.0 - self.rate
        mask = (torch.rand(x.shape, device=x.device, generator=rng) < keep).to(x.dtype)
        return x * mask / keep

    def get_config(self):
        return {"name": self.name, "trainable": True, "rate": self.rate,
                "noise_shape": None, "seed": None}


class Flatten(Layer):
    base_name = "flatten"

    def compute_output_shape(self, s):
        n = 1
        for d in s:
            n *= d
        return (n,)

    def __init__(self, data_format=None, **kw):
        super().__init__(**kw)
        self.data_format = _fmt(data_format)

    def call(self, x, training, rng=None):
        if self.data_format == "channels_first" and x.ndim == 4:
            x = x.permute(0, 2, 3, 1)        # Keras: channels_first inputs are flattened in channels_last order
        return x.reshape(x.shape[0], -1)     # NHWC -> (h, w, c) order, as Keras

    def get_config(self):
        return {"name": self.name, "trainable": True, "data_format": self.data_format}


class Dense(Layer):
    base_name = "dense"

    def __init__(self, units, activation=None, use_bias=True, **kw):
        super().__init__(**kw)
        self.units = int(units)
        self.activation = activation
        self.use_bias = use_bias

    def compute_output_shape(self, s):
        return tuple(s[:-1]) + (self.units,)

    def build(self, input_shape):
        super().build(input_shape)
        fin = input_shape[-1]
        specs = [("kernel", (fin, self.units),
                  lambda g: glorot_uniform((fin, self.units), fin, self.units, g))]
        if self.use_bias:
            specs.append(("bias", (self.units,), lambda g: torch.zeros(self.units)))
        return specs

    def call(self, x, training, rng=None):
        y = x @ self.weights["kernel"]
        if self.use_bias:
            y = y + self.weights["bias"]
        return _ACTIVATIONS[self.activation](y)

    def get_config(self):
        return {"name": self.name, "trainable": True, "units": self.units,
                "activation": self.activation or "linear", "use_bias": self.use_bias,
                "kernel_initializer": {"class_name": "VarianceScaling", "config": {
                    "scale": 1.0, "mode": "fan_avg", "distribution": "uniform", "seed": None}},
                "bias_initializer": {"class_name": "Zeros", "config": {}},
                "kernel_regularizer": None, "bias_regularizer": None,
                "activity_regularizer": None, "kernel_constraint": None, "bias_constraint": None}


class Activation(Layer):
    base_name = "activation"

    def __init__(self, activation, **kw):
        super().__init__(**kw)
        self.activation = activation

    def call(self, x, training, rng=None):
        return _ACTIVATIONS[self.activation](x)

    def get_config(self):
        return {"name": self.name, "trainable": True, "activation": self.activation}


LAYER_CLASSES = {c.__name__: c for c in
                 (InputLayer, Conv2D, MaxPooling2D, AveragePooling2D, GlobalAveragePooling2D,
                  Dropout, Flatten, Dense, Activation)}


def from_config(class_name: str, config: Dict) -> Layer:
    cls = LAYER_CLASSES[class_name]
    c = dict(config)
    kw = {"name": c.get("name")}
    if "batch_input_shape" in c and c["batch_input_shape"] is not None:
        kw["input_shape"] = tuple(c["batch_input_shape"][1:])
    if cls is Conv2D:
        return Conv2D(c["filters"], tuple(c["kernel_size"]), tuple(c.get("strides", (1, 1))),
                      c.get("padding", "valid"),
                      None if c.get("activation") == "linear" else c.get("activation"),
                      c.get("use_bias", True), data_format=c.get("data_format"), **kw)
    if cls in (MaxPooling2D, AveragePooling2D):
        return cls(tuple(c["pool_size"]), tuple(c["strides"]) if c.get("strides") else None,
                   c.get("padding", "valid"), data_format=c.get("data_format"), **kw)
    if cls is Flatten:
        return Flatten(data_format=c.get("data_format"), **kw)
    if cls is Dropout:
        return Dropout(c["rate"], **kw)
    if cls is Dense:
        return Dense(c["units"], None if c.get("activation") == "linear" else c.get("activation"),
                     c.get("use_bias", True), **kw)
    if cls is Activation:
        return Activation(c["activation"], **kw)
    return cls(**kw)
