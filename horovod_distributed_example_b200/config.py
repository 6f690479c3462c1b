"""Environment-variable configuration, read once inside ``hvd.init()``.

The reference has no flag system at all: its whole configuration surface is the
``PS_MODEL_PATH`` env var (reference ``mnist_keras.py:26-27``,
``tensorflow2_keras_mnist.py:21-22``) plus Horovod's ``HOROVOD_*`` variables,
which Horovod reads once at init.  We keep that contract: every knob is an
``HVD_*`` variable, the matching ``HOROVOD_*`` spelling is accepted as an alias,
and nothing is re-read after ``init()``.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Optional


def _env(name: str, default: Optional[str] = None) -> Optional[str]:
    """Look up ``HVD_<name>`` first, then the ``HOROVOD_<name>`` alias."""
    for prefix in ("HVD_", "HOROVOD_"):
        v = os.environ.get(prefix + name)
        if v is not None and v != "":
            return v
    return default


def _env_int(name: str, default: int) -> int:
    v = _env(name)
    if v is None:
        return default
    try:
        return int(float(v))
    except ValueError as e:  # loud: a mistyped knob must not be silently ignored
        raise ValueError(f"HVD_{name}={v!r} is not an integer") from e


def _env_float(name: str, default: float) -> float:
    v = _env(name)
    if v is None:
        return default
    try:
        return float(v)
    except ValueError as e:
        raise ValueError(f"HVD_{name}={v!r} is not a number") from e


def _env_bool(name: str, default: bool = False) -> bool:
    v = _env(name)
    if v is None:
        return default
    return v.strip().lower() not in ("0", "false", "no", "off", "")


ALLREDUCE_VARIANTS = ("auto", "ll", "oneshot", "twoshot", "nvls", "nccl")


@dataclass
class Config:
    #: bytes; buckets are closed when adding the next tensor would cross this.
    #: Horovod's default fusion buffer is 64 MB (HOROVOD_FUSION_THRESHOLD).
    fusion_threshold: int = 64 * 1024 * 1024
    #: which peer-memory allreduce kernel to use; ``auto`` consults the variant table.
    allreduce_variant: str = "auto"
    #: path of a Chrome-trace JSON written by rank 0 (HOROVOD_TIMELINE equivalent).
    timeline: Optional[str] = None
    timeline_mark_cycles: bool = False
    log_level: str = "WARNING"
    #: every N steps compare a parameter checksum across ranks (0 = off).
    debug_checksum: int = 0
    debug_poison: bool = False         # NaN-fill consumed gradient ranges after every update (stale-read trap)
    nvtx: bool = False                 # NVTX ranges around steps / buckets / collectives
    #: "rank:step" - that rank skips its collective at that step (tests the watchdog).
    fault_inject: Optional[str] = None
    #: seconds after which a rank that is still waiting for its peers logs a WARNING (Horovod's
    #: HOROVOD_STALL_CHECK_TIME_SECONDS: a warning, not an abort).
    stall_check_time: float = 60.0
    #: seconds after which an in-kernel barrier gives up, raises the error word and aborts the collective
    #: (HOROVOD_STALL_SHUTDOWN_TIME_SECONDS).  0 = ten times the warning time (600 s by default), so a
    #: rank-0-only pause such as a large checkpoint or an export cannot corrupt the other ranks.
    stall_shutdown_time: float = 0.0
    #: symmetric-memory backend: auto | vmm | ipc
    symm_backend: str = "auto"
    #: where the measured variant table is cached.
    autotune_log: Optional[str] = None
    autotune: bool = False
    #: number of CTAs for the collective kernels (0 = pick from message size).
    comm_ctas: int = 0
    extra: dict = field(default_factory=dict)

    @staticmethod
    def from_env() -> "Config":
        c = Config()
        c.fusion_threshold = _env_int("FUSION_THRESHOLD", c.fusion_threshold)
        c.allreduce_variant = (_env("ALLREDUCE_VARIANT", c.allreduce_variant) or "auto").lower()
        if c.allreduce_variant not in ALLREDUCE_VARIANTS:
            raise ValueError(
                f"HVD_ALLREDUCE_VARIANT={c.allreduce_variant!r}; expected one of {ALLREDUCE_VARIANTS}")
        c.timeline = _env("TIMELINE")
        c.timeline_mark_cycles = _env_bool("TIMELINE_MARK_CYCLES")
        c.log_level = (_env("LOG_LEVEL", c.log_level) or "WARNING").upper()
        c.debug_checksum = _env_int("DEBUG_CHECKSUM", 0)
        c.debug_poison = _env_int("DEBUG_POISON", 0) != 0
        c.nvtx = _env_int("NVTX", 0) != 0
        c.fault_inject = _env("FAULT_INJECT")
        c.stall_check_time = _env_float("STALL_CHECK_TIME_SECONDS", c.stall_check_time)
        c.stall_shutdown_time = _env_float("STALL_SHUTDOWN_TIME_SECONDS", c.stall_shutdown_time)
        c.symm_backend = (_env("SYMM_BACKEND", c.symm_backend) or "auto").lower()
        c.autotune_log = _env("AUTOTUNE_LOG")
        c.autotune = _env_bool("AUTOTUNE")
        c.comm_ctas = _env_int("COMM_CTAS", 0)
        return c

    def barrier_abort_time(self) -> float:
        """Seconds an in-kernel barrier may spin before it aborts (0 = spin forever)."""
        if self.stall_shutdown_time > 0:
            return self.stall_shutdown_time
        return 10.0 * self.stall_check_time if self.stall_check_time > 0 else 0.0

    def parsed_fault(self):
        """Return (rank, step) from HVD_FAULT_INJECT or None."""
        if not self.fault_inject:
            return None
        try:
            r, s = self.fault_inject.split(":")
            return int(r), int(s)
        except Exception as e:
            raise ValueError(f"HVD_FAULT_INJECT={self.fault_inject!r}; expected 'rank:step'") from e


def model_paths():
    """The reference's output-path rule (``mnist_keras.py:26-27``).

    Returns (model_dir, export_dir): ``$PS_MODEL_PATH or ./models`` with
    ``/horovod-mnist`` appended for the model dir.
    """
    base = os.environ.get("PS_MODEL_PATH", os.getcwd() + "/models")
    return os.path.abspath(base + "/horovod-mnist"), os.path.abspath(base)
