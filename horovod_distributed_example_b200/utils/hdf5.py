"""Python face of the native minimal HDF5 writer/reader (``csrc/io/hdf5_min.cpp``).

There is no h5py/libhdf5 on this image, so the framework carries its own implementation
of the HDF5 subset Keras checkpoints use.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Sequence, Union

import numpy as np

from ..ops import native

DT_F32, DT_I64, DT_F64, DT_I32, DT_STR = 0, 1, 2, 3, 4
_NP2DT = {np.dtype("float32"): DT_F32, np.dtype("int64"): DT_I64,
          np.dtype("float64"): DT_F64, np.dtype("int32"): DT_I32}
_DT2NP = {v: k for k, v in _NP2DT.items()}


class H5Error(RuntimeError):
    pass


def _err(lib) -> str:
    return lib.hvd_h5_last_error().decode("utf8", "replace")


class Writer:
    def __init__(self, path: str):
        self._lib = native.core_lib()
        self._h = self._lib.hvd_h5w_create()
        self.path = path
        self._closed = False

    def _chk(self, rc):
        if rc != 0:
            raise H5Error(_err(self._lib))

    def create_group(self, path: str):
        self._chk(self._lib.hvd_h5w_group(self._h, path.encode()))

    def create_dataset(self, path: str, data):
        a = np.ascontiguousarray(data)
        if a.dtype not in _NP2DT:
            if np.issubdtype(a.dtype, np.floating):
                a = a.astype(np.float32)
            elif np.issubdtype(a.dtype, np.integer):
                a = a.astype(np.int64)
            else:
                raise H5Error(f"unsupported dataset dtype {a.dtype}")
        dims = (ctypes.c_int64 * max(a.ndim, 1))(*a.shape)
        self._chk(self._lib.hvd_h5w_dataset(self._h, path.encode(), _NP2DT[a.dtype], a.ndim, dims,
                                            a.ctypes.data_as(ctypes.c_void_p)))

    def set_attr(self, obj: str, name: str, value: Union[bytes, str, Sequence, np.ndarray]):
        """bytes/str -> scalar fixed-length string; list of bytes/str -> 1-D string array;
        numeric array/scalar -> numeric attribute."""
        if isinstance(value, str):
            value = value.encode("utf8")
        if isinstance(value, bytes):
            self._chk(self._lib.hvd_h5w_attr_str(self._h, obj.encode(), name.encode(), value, len(value)))
            return
        if isinstance(value, (list, tuple)) and (len(value) == 0 or isinstance(value[0], (bytes, str))):
            items = [v.encode("utf8") if isinstance(v, str) else v for v in value]
            width = max([len(v) for v in items] + [1])
            packed = b"".join(v.ljust(width, b"\0") for v in items)
            self._chk(self._lib.hvd_h5w_attr_strs(self._h, obj.encode(), name.encode(), packed,
                                                  len(items), width))
            return
        a = np.ascontiguousarray(value)
        if a.dtype not in _NP2DT:
            a = a.astype(np.float64 if np.issubdtype(a.dtype, np.floating) else np.int64)
        dims = (ctypes.c_int64 * max(a.ndim, 1))(*a.shape)
        self._chk(self._lib.hvd_h5w_attr_num(self._h, obj.encode(), name.encode(), _NP2DT[a.dtype],
                                             a.ndim, dims, a.ctypes.data_as(ctypes.c_void_p)))

    def close(self):
        if self._closed:
            return
        try:
            self._chk(self._lib.hvd_h5w_write(self._h, self.path.encode()))
        finally:
            self._lib.hvd_h5w_free(self._h)
            self._closed = True

    def __enter__(self):
        return self

    def __exit__(self, et, ev, tb):
        if et is None:
            self.close()
        elif not self._closed:
            self._lib.hvd_h5w_free(self._h)
            self._closed = True


class Reader:
    def __init__(self, path: str):
        self._lib = native.core_lib()
        self._h = self._lib.hvd_h5r_open(path.encode())
        if not self._h:
            raise H5Error(_err(self._lib))

    def close(self):
        if self._h:
            self._lib.hvd_h5r_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def kind(self, path: str) -> str:
        return {0: "missing", 1: "group", 2: "dataset"}[self._lib.hvd_h5r_kind(self._h, path.encode())]

    def _names(self, fn, path: str) -> List[str]:
        n = fn(self._h, path.encode(), None, 0)
        if n < 0:
            raise H5Error(_err(self._lib))
        buf = ctypes.create_string_buffer(int(n))
        fn(self._h, path.encode(), buf, n)
        s = buf.value.decode("utf8")
        return s.split("\n") if s else []

    def list(self, path: str = "/") -> List[str]:
        return self._names(self._lib.hvd_h5r_list, path)

    def attr_names(self, obj: str = "/") -> List[str]:
        return self._names(self._lib.hvd_h5r_attr_list, obj)

    def dataset(self, path: str) -> np.ndarray:
        dt, rank, es = ctypes.c_int(), ctypes.c_int(), ctypes.c_int64()
        dims = (ctypes.c_int64 * 8)()
        if self._lib.hvd_h5r_dataset_info(self._h, path.encode(), ctypes.byref(dt), ctypes.byref(rank),
                                          dims, ctypes.byref(es)) != 0:
            raise H5Error(_err(self._lib))
        shape = tuple(dims[i] for i in range(rank.value))
        out = np.empty(shape, dtype=_DT2NP[dt.value])
        if self._lib.hvd_h5r_dataset_read(self._h, path.encode(), out.ctypes.data_as(ctypes.c_void_p),
                                          out.nbytes) != 0:
            raise H5Error(_err(self._lib))
        return out

    def attr(self, obj: str, name: str):
        dt, rank, es = ctypes.c_int(), ctypes.c_int(), ctypes.c_int64()
        dims = (ctypes.c_int64 * 8)()
        if self._lib.hvd_h5r_attr_info(self._h, obj.encode(), name.encode(), ctypes.byref(dt),
                                       ctypes.byref(rank), dims, ctypes.byref(es)) != 0:
            raise H5Error(_err(self._lib))
        shape = tuple(dims[i] for i in range(rank.value))
        n = int(np.prod(shape)) if shape else 1
        if dt.value == DT_STR:
            raw = ctypes.create_string_buffer(n * es.value)
            if self._lib.hvd_h5r_attr_read(self._h, obj.encode(), name.encode(), raw, n * es.value) != 0:
                raise H5Error(_err(self._lib))
            b = raw.raw
            items = [b[i * es.value:(i + 1) * es.value].rstrip(b"\0") for i in range(n)]
            return items if shape else items[0]
        out = np.empty(shape, dtype=_DT2NP[dt.value])
        if self._lib.hvd_h5r_attr_read(self._h, obj.encode(), name.encode(),
                                       out.ctypes.data_as(ctypes.c_void_p), out.nbytes) != 0:
            raise H5Error(_err(self._lib))
        return out if shape else out[()]

    def attrs(self, obj: str = "/") -> Dict:
        return {n: self.attr(obj, n) for n in self.attr_names(obj)}
