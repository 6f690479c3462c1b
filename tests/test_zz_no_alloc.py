"""The allocation-free Dense forward (``hvd_dense_forward_noalloc``), exercised through the C ABI.

Some users of this library cannot call cudaMalloc at all - code embedded in another runtime's memory
pool, or a serving loop that must never take the allocator's device-wide lock - so they cannot hand
``hvd_dense_forward`` the split-K scratch buffer it wants.  The no-alloc entry point keeps that buffer
inside the library image instead.  This test checks that it computes the same thing as the ordinary
entry point, twice in a row (the built-in buffer is shared between calls and must be re-cleared),
and that it refuses problems larger than the built-in buffer with a readable error.

The file name sorts last on purpose: if this fails under ``pytest -x`` every other test has already run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device is visible")
    from horovod_distributed_example_b200.ops import native
    lib = native.cuda_lib(required=True)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.hvd_dense_forward.restype = ci
    lib.hvd_dense_forward.argtypes = [vp, vp, vp, vp, ci, ci, ci, ci, ci, vp, vp]
    lib.hvd_dense_forward_noalloc.restype = ci
    lib.hvd_dense_forward_noalloc.argtypes = [vp, vp, vp, vp, ci, ci, ci, ci, ci, vp]
    lib.hvd_dense_forward_noalloc_error.restype = ctypes.c_char_p
    lib.hvd_dense_forward_noalloc_max_elems.restype = ctypes.c_longlong
    return lib


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-12)).item()


def test_noalloc_matches_the_ordinary_entry_point_twice_in_a_row():
    lib = _lib()
    torch.manual_seed(0)
    s = torch.cuda.current_stream().cuda_stream
    # the reference model's Dense(128): x[128, 9216] * W[9216, 128], relu
    shapes = [(128, 128, 9216, 1, 36), (96, 16, 128, 0, 1)]
    for call in range(2):                              # two calls in a row share the built-in buffer
        for M, N, K, relu, ks in shapes:
            x = torch.randn(M, K, device="cuda")
            W = torch.randn(K, N, device="cuda") * 0.05
            b = torch.randn(N, device="cuda")
            want = torch.empty(M, N, device="cuda")
            got = torch.full((M, N), float("nan"), device="cuda")
            scratch = torch.zeros(M * N, device="cuda")
            assert lib.hvd_dense_forward(x.data_ptr(), W.data_ptr(), b.data_ptr(), want.data_ptr(), M, N, K, relu, ks,
                                         scratch.data_ptr(), s) == 0
            rc = lib.hvd_dense_forward_noalloc(x.data_ptr(), W.data_ptr(), b.data_ptr(), got.data_ptr(), M, N, K,
                                               relu, ks, s)
            assert rc == 0, lib.hvd_dense_forward_noalloc_error().decode()
            torch.cuda.synchronize()
            assert not got.isnan().any()
            # same kernels, only the order of the split-K atomic adds differs: TF32-level tolerance
            assert _rel(got, want) < 3e-3, (call, M, N, K)
            ref = x.double() @ W.double() + b.double()
            ref = torch.relu(ref) if relu else ref
            assert _rel(got, ref) < 3e-3


def test_noalloc_rejects_sizes_beyond_the_builtin_buffer():
    lib = _lib()
    cap = lib.hvd_dense_forward_noalloc_max_elems()
    assert cap == 256 * 1024
    M, N, K = 1024, 512, 32                            # M*N = 524288 > cap
    x = torch.zeros(M, K, device="cuda")
    W = torch.zeros(K, N, device="cuda")
    y = torch.full((M, N), 7.0, device="cuda")
    rc = lib.hvd_dense_forward_noalloc(x.data_ptr(), W.data_ptr(), None, y.data_ptr(), M, N, K, 0, 1,
                                       torch.cuda.current_stream().cuda_stream)
    assert rc == -100
    msg = lib.hvd_dense_forward_noalloc_error().decode()
    assert "exceed" in msg and "262144" in msg
    torch.cuda.synchronize()
    assert (y == 7.0).all()                            # nothing was launched
