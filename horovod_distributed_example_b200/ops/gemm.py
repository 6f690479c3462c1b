"""Python binding of the tcgen05 TF32 GEMM (``csrc/kernels/tc_gemm.cu``)."""
from __future__ import annotations

import ctypes

import torch

from . import native

EPI_STORE, EPI_ATOMIC_ADD, EPI_BIAS_RELU, EPI_UNPOOL, EPI_RS_PUSH = 0, 1, 2, 3, 4


def _lib():
    lib = native.cuda_lib()
    if not getattr(lib, "_hvd_gemm_declared", False):
        lib.hvd_tc_gemm.restype = ctypes.c_int
        lib.hvd_tc_gemm.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p,
                                    ctypes.c_uint, ctypes.c_uint, ctypes.c_uint, ctypes.c_uint,
                                    ctypes.c_void_p, ctypes.c_float]
        lib.hvd_tc_gemm_ex.restype = ctypes.c_int
        lib.hvd_tc_gemm_ex.argtypes = list(lib.hvd_tc_gemm.argtypes) + [ctypes.c_void_p, ctypes.c_void_p]
        lib._hvd_gemm_declared = True
    return lib


def tc_gemm(A: torch.Tensor, B: torch.Tensor, M: int, N: int, K: int, a_mn: bool = False, b_mn: bool = False,
            bn: int = 128, k_split: int = 1, epilogue: int = EPI_STORE, bias: torch.Tensor = None,
            out: torch.Tensor = None, dbg=(0, 0, 0, 0), mask: torch.Tensor = None, scale: float = 1.0,
            stream=None, rs=None, wait=None) -> torch.Tensor:
    """D[M,N] = A*B on the tensor cores (TF32 in, fp32 accumulate).

    ``A`` is ``[M,K]`` (``a_mn=False``) or ``[K,M]`` (``a_mn=True``); ``B`` is ``[N,K]`` or ``[K,N]``.
    With ``k_split>1`` pass ``epilogue=EPI_ATOMIC_ADD`` and a zeroed ``out``.
    ``epilogue=EPI_RS_PUSH`` with ``rs`` (a ``parallel.pipeline.HvdRsPush``) sends each 128x128 output tile to
    the rank that owns it instead of writing ``out`` (which may then be None); ``wait`` (``HvdParamWait``) makes
    the kernel wait for peer-delivered weights before its first load."""
    assert A.is_cuda and B.is_cuda and A.dtype == torch.float32 and B.dtype == torch.float32
    assert A.is_contiguous() and B.is_contiguous()
    if epilogue == EPI_RS_PUSH or rs is not None or wait is not None:
        rc = _lib().hvd_tc_gemm_ex(A.data_ptr(), B.data_ptr(), out.data_ptr() if out is not None else None, M, N, K,
                                   int(a_mn), int(b_mn), bn, k_split, epilogue,
                                   bias.data_ptr() if bias is not None else None,
                                   out.stride(0) if out is not None else N,
                                   stream if stream is not None else torch.cuda.current_stream().cuda_stream, *dbg,
                                   mask.data_ptr() if mask is not None else None, float(scale),
                                   ctypes.addressof(rs) if rs is not None else None,
                                   ctypes.addressof(wait) if wait is not None else None)
        if rc != 0:
            raise RuntimeError(f"hvd_tc_gemm_ex failed: {rc}")
        return out
    if out is None:
        out = torch.zeros(M, N, device=A.device, dtype=torch.float32) if epilogue == EPI_ATOMIC_ADD \
            else torch.empty(M, N, device=A.device, dtype=torch.float32)
    rc = _lib().hvd_tc_gemm(A.data_ptr(), B.data_ptr(), out.data_ptr(), M, N, K, int(a_mn), int(b_mn), bn, k_split,
                            epilogue, bias.data_ptr() if bias is not None else None, out.stride(0),
                            stream if stream is not None else torch.cuda.current_stream().cuda_stream, *dbg,
                            mask.data_ptr() if mask is not None else None, float(scale))
    if rc != 0:
        raise RuntimeError(f"hvd_tc_gemm failed: {rc}")
    return out
