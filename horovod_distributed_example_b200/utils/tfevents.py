"""TensorBoard event-file writer/reader over the native TFRecord encoder
(``csrc/core/core.cpp``: masked CRC32C framing + hand-encoded Event/Summary protobufs).

Stands in for the event files Keras' ``TensorBoard`` callback writes in the reference
(``mnist_keras.py:105``, ``tensorflow2_keras_mnist.py:89``).  The reader exists for tests
and for the metrics sink; it parses only what the writer emits (scalar summaries).
"""
from __future__ import annotations

import os
import socket
import struct
import time
from typing import Iterator, List, Tuple

from ..ops import native


class EventWriter:
    def __init__(self, log_dir: str):
        os.makedirs(log_dir, exist_ok=True)
        self.path = os.path.join(
            log_dir, f"events.out.tfevents.{int(time.time())}.{socket.gethostname()}.{os.getpid()}.hvd")
        self._lib = native.core_lib()
        self._h = self._lib.hvd_ev_open(self.path.encode())
        if not self._h:
            raise OSError(f"cannot open event file {self.path}")

    def add_scalar(self, tag: str, value: float, step: int, wall_time: float = 0.0) -> None:
        if self._h:
            self._lib.hvd_ev_scalar(self._h, tag.encode(), float(value), int(step), float(wall_time))

    def flush(self) -> None:
        if self._h:
            self._lib.hvd_ev_flush(self._h)

    def close(self) -> None:
        if self._h:
            self._lib.hvd_ev_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _masked_crc(data: bytes) -> int:
    lib = native.core_lib()
    c = lib.hvd_crc32c(data, len(data))
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def _varint(b: bytes, i: int) -> Tuple[int, int]:
    v = s = 0
    while True:
        x = b[i]
        i += 1
        v |= (x & 0x7F) << s
        s += 7
        if x < 0x80:
            return v, i


def _fields(b: bytes) -> Iterator[Tuple[int, int, object]]:
    i = 0
    while i < len(b):
        key, i = _varint(b, i)
        f, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _varint(b, i)
        elif wt == 1:
            v = b[i:i + 8]
            i += 8
        elif wt == 2:
            n, i = _varint(b, i)
            v = b[i:i + n]
            i += n
        elif wt == 5:
            v = b[i:i + 4]
            i += 4
        else:
            raise ValueError("bad wire type")
        yield f, wt, v


def read_scalars(path: str) -> List[Tuple[str, int, float]]:
    """Return [(tag, step, value)] from an event file, verifying both CRCs of each record."""
    out = []
    with open(path, "rb") as f:
        data = f.read()
    i = 0
    while i + 12 <= len(data):
        hdr = data[i:i + 8]
        (n,) = struct.unpack("<Q", hdr)
        (c1,) = struct.unpack("<I", data[i + 8:i + 12])
        if c1 != _masked_crc(hdr):
            raise ValueError("corrupt record length CRC")
        body = data[i + 12:i + 12 + n]
        (c2,) = struct.unpack("<I", data[i + 12 + n:i + 16 + n])
        if c2 != _masked_crc(body):
            raise ValueError("corrupt record data CRC")
        i += 16 + n
        step, summary = 0, None
        for fno, _wt, v in _fields(body):
            if fno == 2:
                step = v
            elif fno == 5:
                summary = v
        if summary is None:
            continue
        for fno, _wt, val in _fields(summary):
            if fno != 1:
                continue
            tag, sv = None, None
            for f2, _w2, v2 in _fields(val):
                if f2 == 1:
                    tag = v2.decode()
                elif f2 == 2:
                    (sv,) = struct.unpack("<f", v2)
            if tag is not None and sv is not None:
                out.append((tag, step, sv))
    return out
