"""NVTX ranges and the native log level.

``HVD_NVTX=1`` wraps every training step, every gradient bucket and every explicit collective in an NVTX
range so an Nsight timeline shows them by name next to the kernels.  The ranges are emitted by the native
core (``csrc/core/diag.cpp``, header-only NVTX3), which costs a branch when no profiler is attached; they
are still opt-in because a ctypes call per range is ~1 us of host time in a ~100 us step.

``configure`` also forwards the rank and ``HVD_LOG_LEVEL`` to the native logger so C++ messages carry the
same ``[hvd <rank>]`` prefix as the Python ones.
"""
from __future__ import annotations

import contextlib

_enabled = False
_lib = None


def configure(enabled: bool, rank: int = -1, level: str = "WARNING") -> None:
    global _enabled, _lib
    from ..ops import native
    lib = native.core_lib(required=False)
    if lib is None:
        _enabled = False
        return
    lib.hvd_log_set_rank(int(rank))
    lib.hvd_log_set_level(lib.hvd_log_parse_level(level.encode()))
    _lib = lib
    _enabled = bool(enabled)


def enabled() -> bool:
    return _enabled


def push(name: str) -> None:
    if _enabled:
        _lib.hvd_nvtx_push(name.encode())


def pop() -> None:
    if _enabled:
        _lib.hvd_nvtx_pop()


def mark(name: str) -> None:
    if _enabled:
        _lib.hvd_nvtx_mark(name.encode())


@contextlib.contextmanager
def range(name: str):
    if not _enabled:
        yield
        return
    _lib.hvd_nvtx_push(name.encode())
    try:
        yield
    finally:
        _lib.hvd_nvtx_pop()


def native_log(level: int, msg: str) -> None:
    """Emit through the C++ logger (tests use it to check level filtering)."""
    if _lib is not None:
        _lib.hvd_log(int(level), msg.encode())
