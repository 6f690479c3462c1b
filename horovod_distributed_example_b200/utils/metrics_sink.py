"""Local metrics sink: the stand-in for ``gradient_utils.metrics.init(sync_tensorboard=True)``
(reference ``mnist_keras.py:22-23``), which forwards TensorBoard scalars to the Gradient
platform where the CI check ``tensorflow:loss in 0.0..0.3 (mean)`` runs
(``.ps_project/config.yaml:8-11``).

There is no Gradient backend here: scalars recorded by the TensorBoard callback are
appended to a JSONL file, and ``check()`` evaluates a GradientCI-style range check on
them so the reference's only acceptance test can run in-repo.
"""
from __future__ import annotations

import json
import os
import time
from typing import Dict, List, Optional

_state = {"enabled": False, "path": None, "fh": None, "values": {}}


def init(sync_tensorboard: bool = True, path: Optional[str] = None) -> None:
    """Enable the sink.  ``path`` defaults to ``$HVD_METRICS_PATH`` or ``./metrics.jsonl``."""
    _state["enabled"] = bool(sync_tensorboard)
    _state["path"] = path or os.environ.get("HVD_METRICS_PATH", os.path.join(os.getcwd(), "metrics.jsonl"))
    _state["values"] = {}
    if _state["fh"] is not None:
        _state["fh"].close()
        _state["fh"] = None


def record(tag: str, value: float, step: int) -> None:
    if not _state["enabled"]:
        return
    _state["values"].setdefault(tag, []).append(float(value))
    if _state["fh"] is None:
        d = os.path.dirname(os.path.abspath(_state["path"]))
        os.makedirs(d, exist_ok=True)
        _state["fh"] = open(_state["path"], "a")
    _state["fh"].write(json.dumps({"tag": tag, "value": float(value), "step": int(step), "t": time.time()}) + "\n")
    _state["fh"].flush()


def values(tag: str) -> List[float]:
    return list(_state["values"].get(tag, []))


_ALIASES = {"tensorflow:loss": ("epoch_loss", "batch_loss", "loss")}


def check(metric: str, target: str, aggregate: str = "mean") -> Dict:
    """GradientCI-style check, e.g. ``check('tensorflow:loss', '0.0..0.3', 'mean')``."""
    lo, hi = (float(x) for x in target.split(".."))
    vals: List[float] = []
    for t in _ALIASES.get(metric, (metric,)):
        vals = values(t)
        if vals:
            break
    if not vals:
        return {"metric": metric, "passed": False, "reason": "no values recorded"}
    agg = {"mean": sum(vals) / len(vals), "min": min(vals), "max": max(vals), "last": vals[-1]}[aggregate]
    return {"metric": metric, "aggregate": aggregate, "value": agg, "target": target,
            "passed": lo <= agg <= hi}


def close() -> None:
    if _state["fh"] is not None:
        _state["fh"].close()
        _state["fh"] = None
