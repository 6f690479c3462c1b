"""``HVD_TIMELINE=<file>``: Chrome-trace JSON of collective activity, written by rank 0
through the native writer (``csrc/core/core.cpp``).  Equivalent of ``HOROVOD_TIMELINE``
(SURVEY.md section 5.1).  Host ``begin/end`` pairs bracket each collective; the fused
device path adds ``complete`` events with CUDA-event durations per bucket.
"""
from __future__ import annotations

from ..ops import native


class Timeline:
    def __init__(self, path: str, mark_cycles: bool = False):
        self._lib = native.core_lib()
        self._h = self._lib.hvd_tl_open(path.encode())
        if not self._h:
            raise OSError(f"cannot open timeline file {path}")
        self.mark_cycles = mark_cycles
        self.path = path

    def begin(self, tensor: str, activity: str) -> None:
        if self._h:
            self._lib.hvd_tl_event(self._h, tensor.encode(), activity.encode(), b"B", -1, 0)

    def end(self, tensor: str, activity: str) -> None:
        if self._h:
            self._lib.hvd_tl_event(self._h, tensor.encode(), activity.encode(), b"E", -1, 0)

    def complete(self, tensor: str, activity: str, ts_us: int, dur_us: int) -> None:
        if self._h:
            self._lib.hvd_tl_event(self._h, tensor.encode(), activity.encode(), b"X", int(ts_us), int(dur_us))

    def mark_cycle(self) -> None:
        if self._h and self.mark_cycles:
            self._lib.hvd_tl_event(self._h, b"cycle", b"CYCLE_START", b"i", -1, 0)

    def now_us(self) -> int:
        return int(self._lib.hvd_tl_now_us(self._h)) if self._h else 0

    def close(self) -> None:
        if self._h:
            self._lib.hvd_tl_close(self._h)
            self._h = None
