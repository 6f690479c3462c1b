"""Serving export - the stand-in for the reference's TF ``SavedModelBuilder`` block
(``mnist_keras.py:126-140``): a directory ``<export_dir>/<timestamp>/`` holding the
inference model and a predict signature mapping ``input`` -> ``prob`` under the default
serving key with the SERVING tag.

There is no TensorFlow here, so the artefact is: ``model.h5`` (Keras-layout weights +
config, loadable by ``load_model``), ``signature.json`` (the signature-def equivalent) and
``model.ts`` (a TorchScript trace of the inference forward with dropout off, runnable
without this package).
"""
from __future__ import annotations

import json
import os

import torch

SERVING = "serve"
DEFAULT_SERVING_SIGNATURE_DEF_KEY = "serving_default"


def export(model, export_path: str, inputs=None, outputs=None, tags=(SERVING,), torchscript=True) -> str:
    os.makedirs(export_path, exist_ok=False)
    model.save(os.path.join(export_path, "model.h5"), include_optimizer=False)
    in_name = next(iter(inputs or {"input": None}))
    out_name = next(iter(outputs or {"prob": None}))
    sig = {
        "tags": list(tags),
        "signature_def": {DEFAULT_SERVING_SIGNATURE_DEF_KEY: {
            "method_name": "tensorflow/serving/predict",
            "inputs": {in_name: {"dtype": "float32", "shape": [None] + list(model.input_shape)}},
            "outputs": {out_name: {"dtype": "float32", "shape": [None] + list(model.output_shape)}},
        }},
        "clear_devices": True,
    }
    with open(os.path.join(export_path, "signature.json"), "w") as f:
        json.dump(sig, f, indent=1)
    if torchscript:
        cpu_model = _cpu_copy(model)
        example = torch.zeros((1,) + tuple(model.input_shape))
        with torch.no_grad():
            traced = torch.jit.trace(lambda x: cpu_model._forward(x, False), example, check_trace=False)
        traced.save(os.path.join(export_path, "model.ts"))
    return export_path


def _cpu_copy(model):
    """A CPU twin of ``model`` (weights copied) for tracing."""
    from ..keras.models import Sequential
    twin = Sequential.from_config(model.get_config())
    import horovod_distributed_example_b200.keras.models as mm
    saved = mm._default_device
    mm._default_device = lambda: torch.device("cpu")
    try:
        twin.build(model.input_shape)
    finally:
        mm._default_device = saved
    twin.set_weights(model.get_weights())
    for lay in twin.layers:
        for k in list(lay.weights):
            lay.weights[k] = lay.weights[k].detach()
    return twin
