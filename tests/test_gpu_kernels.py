"""Numerics of every sm_100a model kernel against a plain PyTorch fp32/fp64 reference of the same op.
TF32 tensor-core paths: ~1e-3 relative; CUDA-core / elementwise paths: ~1e-5."""
import ctypes
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
B = 128


def _lib():
    if not torch.cuda.is_available():
        pytest.fail("marked gpu but no CUDA device is visible")
    from horovod_distributed_example_b200.models import mnist_cnn
    from horovod_distributed_example_b200.ops import native
    lib = native.cuda_lib(required=True)
    mnist_cnn._declare(lib)
    return lib


def _s():
    return torch.cuda.current_stream().cuda_stream


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-12)).item()


def _state(seed=7):
    st = torch.zeros(8, dtype=torch.int32, device="cuda")
    st[5] = seed
    return st


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_tc_gemm_all_majors(a_mn, b_mn):
    _lib()
    from horovod_distributed_example_b200.ops.gemm import tc_gemm
    torch.manual_seed(1)
    M, N, K = 384, 200, 96 + 32 * a_mn
    A, Bm = torch.randn(M, K, device="cuda"), torch.randn(K, N, device="cuda")
    out = tc_gemm(A.t().contiguous() if a_mn else A, Bm if b_mn else Bm.t().contiguous(), M, N, K, a_mn, b_mn, bn=128)
    assert _rel(out, A.double() @ Bm.double()) < 3e-3


def test_tc_gemm_splitk_and_ktail():
    _lib()
    from horovod_distributed_example_b200.ops.gemm import EPI_ATOMIC_ADD, tc_gemm
    torch.manual_seed(2)
    A, W = torch.randn(128, 9216, device="cuda"), torch.randn(9216, 128, device="cuda")
    out = tc_gemm(A, W, 128, 128, 9216, False, True, 128, k_split=96, epilogue=EPI_ATOMIC_ADD)
    assert _rel(out, A.double() @ W.double()) < 3e-3
    X, G = torch.randn(44, 640, device="cuda"), torch.randn(44, 128, device="cuda")      # K = 44: TMA zero-fills the tail
    dw = tc_gemm(X, G, 640, 128, 44, True, True, 128)
    assert _rel(dw, X.double().t() @ G.double()) < 3e-3


def test_conv1_fwd():
    lib = _lib()
    torch.manual_seed(3)
    x = torch.rand(B, 28, 28, device="cuda")
    w = torch.randn(3, 3, 1, 32, device="cuda") * 0.3
    b = torch.randn(32, device="cuda") * 0.1
    a1 = torch.empty(B, 26, 26, 32, device="cuda")
    z1, z2, st = torch.ones(64, device="cuda"), torch.ones(1000, device="cuda"), _state()
    assert lib.hvd_conv1_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), a1.data_ptr(), B, z1.data_ptr(), 64,
                             z2.data_ptr(), 1000, st.data_ptr(), 1, _s()) == 0
    assert (z1 == 0).all() and (z2 == 0).all() and st[0].item() == 1
    ref = F.relu(F.conv2d(x[:, None], w.permute(3, 2, 0, 1), b)).permute(0, 2, 3, 1)
    torch.testing.assert_close(a1, ref, rtol=1e-5, atol=1e-5)


def _conv2_ref(a1, w2, b2):
    y = F.conv2d(a1.permute(0, 3, 1, 2).double(), w2.permute(3, 2, 0, 1).double(), b2.double())
    return F.relu(y)           # [B,64,24,24]


@pytest.mark.parametrize("batch", [128, 37])
def test_conv2_fprop_eval_and_masks(batch):
    lib = _lib()
    torch.manual_seed(4)
    a1 = torch.relu(torch.randn(batch, 26, 26, 32, device="cuda"))
    w2 = torch.randn(3, 3, 32, 64, device="cuda") * 0.05
    b2 = torch.randn(64, device="cuda") * 0.1
    p2 = torch.full((batch, 12, 12, 64), -7.0, device="cuda")
    mask = torch.zeros(batch, 12, 12, 64, dtype=torch.uint8, device="cuda")
    st = _state()
    assert lib.hvd_conv2_fprop(a1.data_ptr(), w2.data_ptr(), b2.data_ptr(), p2.data_ptr(), mask.data_ptr(),
                               st.data_ptr(), batch, 0, 0.75, 0, _s()) == 0
    y = _conv2_ref(a1, w2, b2)
    pooled, idx = F.max_pool2d(y, 2, return_indices=True)
    ref = pooled.permute(0, 2, 3, 1).float()
    assert _rel(p2, ref) < 3e-3
    # arg-max position: decode and compare where the window max is unambiguous
    iy, ix = (idx // 24) % 2, (idx % 24) % 2
    pos_ref = (iy * 2 + ix).permute(0, 2, 3, 1)
    srt = y.unfold(2, 2, 2).unfold(3, 2, 2).reshape(batch, 64, 12, 12, 4).sort(-1).values
    clear = ((srt[..., 3] - srt[..., 2]) > 1e-2).permute(0, 2, 3, 1)
    assert ((mask & 3).long()[clear] == pos_ref[clear]).all()
    live = (mask & 4) != 0
    assert (live == (p2 > 0)).all()
    # training mode: kept elements are scaled by 1/keep, ~75% kept, mask differs per step
    p2t = torch.empty_like(p2)
    maskt = torch.empty_like(mask)
    assert lib.hvd_conv2_fprop(a1.data_ptr(), w2.data_ptr(), b2.data_ptr(), p2t.data_ptr(), maskt.data_ptr(),
                               st.data_ptr(), batch, 1, 0.75, 0, _s()) == 0
    kept = p2t != 0
    torch.testing.assert_close(p2t[kept], (p2 / 0.75)[kept], rtol=1e-6, atol=1e-6)
    frac = kept.float().sum() / (p2 != 0).float().sum()
    assert 0.74 < frac.item() < 0.76


def test_head_forward_backward():
    lib = _lib()
    torch.manual_seed(5)
    h_pre = torch.randn(B, 128, device="cuda")
    b1 = torch.randn(128, device="cuda") * 0.1
    w2 = (torch.randn(128, 10, device="cuda") * 0.2).requires_grad_()
    b2 = (torch.randn(10, device="cuda") * 0.1).requires_grad_()
    y = torch.randint(0, 10, (B,), device="cuda")
    pre = (h_pre + b1).requires_grad_()
    logits = F.relu(pre) @ w2 + b2
    loss = F.cross_entropy(logits, y)
    loss.backward()
    hp = h_pre.clone()
    dh = torch.empty(B, 128, device="cuda")
    gw, gb, gb1 = torch.zeros(128, 10, device="cuda"), torch.zeros(10, device="cuda"), torch.zeros(128, device="cuda")
    probs = torch.empty(B, 10, device="cuda")
    st = _state()
    assert lib.hvd_head(hp.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), y.data_ptr(), dh.data_ptr(),
                        gw.data_ptr(), gb.data_ptr(), gb1.data_ptr(), probs.data_ptr(), st.data_ptr(), B, 0, 1, 0.5,
                        _s()) == 0
    stf = st.view(torch.float32)
    assert abs(stf[3].item() - loss.item()) < 1e-5
    assert abs(stf[4].item() - (logits.argmax(1) == y).float().mean().item()) < 1e-6     # training: fraction
    assert (hp == 0).all()                                       # split-K accumulator re-armed
    torch.testing.assert_close(probs, F.softmax(logits, 1).detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(gw, w2.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(gb, b2.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(dh, pre.grad, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(gb1, pre.grad.sum(0), rtol=1e-4, atol=1e-6)


def test_dense1_dx_unpool_epilogue():
    lib = _lib()
    from horovod_distributed_example_b200.ops.gemm import EPI_UNPOOL, tc_gemm
    torch.manual_seed(6)
    a1 = torch.relu(torch.randn(B, 26, 26, 32, device="cuda"))
    w2 = torch.randn(3, 3, 32, 64, device="cuda") * 0.05
    b2 = torch.randn(64, device="cuda") * 0.1
    p2 = torch.empty(B, 9216, device="cuda")
    mask = torch.empty(B, 9216, dtype=torch.uint8, device="cuda")
    assert lib.hvd_conv2_fprop(a1.data_ptr(), w2.data_ptr(), b2.data_ptr(), p2.data_ptr(), mask.data_ptr(),
                               _state().data_ptr(), B, 0, 1.0, 0, _s()) == 0
    dh = torch.randn(B, 128, device="cuda")
    W1 = torch.randn(9216, 128, device="cuda") * 0.05
    dy2 = torch.full((B, 24, 24, 64), 9.0, device="cuda")
    tc_gemm(dh, W1, B, 9216, 128, False, False, 128, epilogue=EPI_UNPOOL, out=dy2.view(-1, 64), mask=mask, scale=1.0)
    # reference: gradient wrt flatten routed back through maxpool+relu using OUR arg-max (ties aside)
    dp = (dh.double() @ W1.double().t()).view(B, 12, 12, 64)
    pos = (mask & 3).view(B, 12, 12, 64).long()
    live = ((mask & 4) != 0).view(B, 12, 12, 64)
    ref = torch.zeros(B, 24, 24, 64, dtype=torch.float64, device="cuda")
    for p_ in range(4):
        sel = (pos == p_) & live
        ref[:, p_ // 2::2, p_ % 2::2, :] = torch.where(sel, dp, torch.zeros_like(dp))
    assert _rel(dy2, ref) < 3e-3
    assert ((dy2 != 0).sum(dim=(1, 2, 3)) <= 12 * 12 * 64).all()


def test_conv2_dgrad_and_wgrad_and_conv1_wgrad():
    lib = _lib()
    torch.manual_seed(7)
    x = torch.rand(B, 28, 28, device="cuda")
    a1 = torch.randn(B, 26, 26, 32, device="cuda")               # pre-"relu" sign pattern matters for the mask
    a1r = torch.relu(a1)
    w2 = torch.randn(3, 3, 32, 64, device="cuda") * 0.05
    dy2 = torch.randn(B, 24, 24, 64, device="cuda") * (torch.rand(B, 24, 24, 64, device="cuda") < 0.25)
    # reference via autograd
    a_in = a1r.permute(0, 3, 1, 2).double().requires_grad_()
    wt = w2.permute(3, 2, 0, 1).double().requires_grad_()
    y = F.conv2d(a_in, wt)
    y.backward(dy2.permute(0, 3, 1, 2).double())
    da_ref = (a_in.grad.permute(0, 2, 3, 1) * (a1r > 0)).float()
    dw_ref = wt.grad.permute(2, 3, 1, 0).float()                  # OIHW -> HWIO
    da1 = torch.full_like(a1, 5.0)
    f1w = torch.zeros(3, 3, 1, 32, device="cuda")
    f1b = torch.zeros(32, device="cuda")
    lib.hvd_conv2_dgrad.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 4
    assert lib.hvd_conv2_dgrad(dy2.data_ptr(), w2.data_ptr(), a1r.data_ptr(), da1.data_ptr(), B, 0, _s(),
                               x.data_ptr(), f1w.data_ptr(), f1b.data_ptr()) == 0
    assert _rel(da1, da_ref) < 3e-3
    for ctas in (5, 3):            # many tiles per CTA: both epilogue warp sets, staging/TMEM stage reuse, odd tails
        da1b = torch.full_like(a1, 5.0)
        f1w2 = torch.zeros(3, 3, 1, 32, device="cuda")
        f1b2 = torch.zeros(32, device="cuda")
        assert lib.hvd_conv2_dgrad(dy2.data_ptr(), w2.data_ptr(), a1r.data_ptr(), da1b.data_ptr(), B, ctas, _s(),
                                   x.data_ptr(), f1w2.data_ptr(), f1b2.data_ptr()) == 0
        assert _rel(da1b, da_ref) < 3e-3
        assert _rel(f1w2, f1w) < 1e-3 and _rel(f1b2, f1b) < 1e-3
    gw = torch.zeros(3, 3, 32, 64, device="cuda")
    gb = torch.zeros(64, device="cuda")
    assert lib.hvd_conv2_wgrad(a1r.data_ptr(), dy2.data_ptr(), gw.data_ptr(), gb.data_ptr(), B, 0, _s()) == 0
    assert _rel(gw, dw_ref) < 3e-3
    assert _rel(gb, dy2.double().sum(dim=(0, 1, 2))) < 1e-4          # bias gradient summed inside the wgrad kernel
    for ctas in (48, 148, 7):                                        # uneven strip ranges, accumulate semantics
        gw2 = torch.ones(3, 3, 32, 64, device="cuda")
        gb2 = torch.ones(64, device="cuda")
        assert lib.hvd_conv2_wgrad(a1r.data_ptr(), dy2.data_ptr(), gw2.data_ptr(), gb2.data_ptr(), B, ctas, _s()) == 0
        assert _rel(gw2 - 1.0, dw_ref) < 3e-3 and _rel(gb2 - 1.0, dy2.double().sum(dim=(0, 1, 2))) < 1e-3
    # wgrad + dgrad (+ conv1 wgrad) as one kernel
    lib.hvd_conv2_bwd.argtypes = [ctypes.c_void_p] * 8 + [ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_int] * 2
    for ctas, nw, t1 in ((0, 0, -1), (37, 5, 0), (148, 64, 3), (2, 1, 7)):
        gw4, gb4 = torch.zeros(3, 3, 32, 64, device="cuda"), torch.zeros(64, device="cuda")
        f1w4, f1b4 = torch.zeros(3, 3, 1, 32, device="cuda"), torch.zeros(32, device="cuda")
        assert lib.hvd_conv2_bwd(dy2.data_ptr(), w2.data_ptr(), a1r.data_ptr(), x.data_ptr(), gw4.data_ptr(), gb4.data_ptr(),
                                 f1w4.data_ptr(), f1b4.data_ptr(), B, ctas, _s(), None, None, nw, t1) == 0
        assert _rel(gw4, dw_ref) < 3e-3 and _rel(gb4, dy2.double().sum(dim=(0, 1, 2))) < 1e-3
        assert _rel(f1w4, f1w) < 1e-3 and _rel(f1b4, f1b) < 1e-3
    gb3 = torch.zeros(64, device="cuda")
    lib.hvd_colsum64.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p]
    assert lib.hvd_colsum64(dy2.data_ptr(), gb3.data_ptr(), B * 576, _s()) == 0
    assert _rel(gb3, dy2.double().sum(dim=(0, 1, 2))) < 1e-4
    # conv1 wgrad (CUDA cores)
    g1w = torch.zeros(3, 3, 1, 32, device="cuda")
    g1b = torch.zeros(32, device="cuda")
    assert lib.hvd_conv1_wgrad(x.data_ptr(), da_ref.contiguous().data_ptr(), g1w.data_ptr(), g1b.data_ptr(), B, _s()) == 0
    xin = x[:, None].double().requires_grad_()
    w1 = torch.zeros(32, 1, 3, 3, dtype=torch.float64, device="cuda", requires_grad=True)
    F.conv2d(xin, w1).backward(da_ref.permute(0, 3, 1, 2).double())
    assert _rel(g1w, w1.grad.permute(2, 3, 1, 0)) < 1e-4
    assert _rel(g1b, da_ref.double().sum(dim=(0, 1, 2))) < 1e-4
    # the same gradients from the tensor-core GEMM fused into the dgrad kernel (TF32 operands)
    assert _rel(f1w, w1.grad.permute(2, 3, 1, 0)) < 3e-3
    assert _rel(f1b, da_ref.double().sum(dim=(0, 1, 2))) < 3e-3


def _models(rate1, rate2, opt_name):
    import horovod_distributed_example_b200 as hvd
    from horovod_distributed_example_b200 import keras
    hvd.init()

    def build():
        keras.backend.clear_session()
        m = keras.Sequential([
            keras.layers.Conv2D(32, (3, 3), activation='relu', input_shape=(28, 28, 1)),
            keras.layers.Conv2D(64, (3, 3), activation='relu'),
            keras.layers.MaxPooling2D((2, 2)), keras.layers.Dropout(rate1), keras.layers.Flatten(),
            keras.layers.Dense(128, activation='relu'), keras.layers.Dropout(rate2),
            keras.layers.Dense(10, activation='softmax')], seed=11)
        o = keras.optimizers.Adadelta(1.0) if opt_name == "adadelta" else keras.optimizers.Adam(0.001)
        m.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=hvd.DistributedOptimizer(o),
                  metrics=['accuracy'])
        return m
    return hvd, build


@pytest.mark.parametrize("opt_name,env", [("adadelta", {}), ("adam", {}),
                                          ("adadelta", {"HVD_BUCKET_MODE": "overlap"}),        # bucket 0 on a side stream
                                          ("adam", {"HVD_FUSED_BWD": "1"}),                     # wgrad + dgrad as one kernel
                                          ("adadelta", {"HVD_FUSE_HEAD": "1"}),                 # dense1 fwd + head as one kernel
                                          ("adadelta", {"HVD_BUCKET_MODE": "overlap", "HVD_RESERVE_SMS": "0"})])
def test_engine_matches_torch_oracle_without_dropout(opt_name, env, monkeypatch):
    """Whole fused step (CUDA graph, fused update) vs the torch-autograd oracle, dropout rate 0 - in the default
    configuration and in the selectable step shapes."""
    _lib()
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    hvd, build = _models(0.0, 0.0, opt_name)
    rng = np.random.default_rng(0)
    X = rng.random((3, B, 28, 28, 1), dtype=np.float32)
    Y = rng.integers(0, 10, (3, B))
    m = build()
    m._ensure_built((B, 28, 28, 1))
    assert type(m._backend).__name__ == "MnistEngine", "native engine not selected on a B200"
    monkeypatch.setenv("HVD_ENGINE", "torch")
    ref = build()
    ref._ensure_built((B, 28, 28, 1))
    assert type(ref._backend).__name__ == "_TorchBackend"
    ref.set_weights(m.get_weights())
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    for i in range(3):
        l1 = m.train_on_batch(X[i], Y[i])
        l2 = ref.train_on_batch(X[i], Y[i])
        assert abs(l1[0] - l2[0]) < 5e-3 * max(1.0, abs(l2[0])), (i, l1, l2)
        assert abs(l1[1] - l2[1]) <= 2.0 / B
    for a, b in zip(m.get_weights(), ref.get_weights()):
        assert np.abs(a - b).max() < 2e-3 * max(1.0, np.abs(b).max()) + (1e-2 if opt_name == "adadelta" else 2e-3)
    ev1, ev2 = m.evaluate(X[0], Y[0], batch_size=B, verbose=0), ref.evaluate(X[0], Y[0], batch_size=B, verbose=0)
    assert abs(ev1[0] - ev2[0]) < 5e-3 and abs(ev1[1] - ev2[1]) <= 2.0 / B
    np.testing.assert_allclose(m.predict(X[1][:50], batch_size=50), ref.predict(X[1][:50], batch_size=50), atol=5e-3)
    hvd.shutdown()


def test_engine_timeline_has_one_update_event_per_step(tmp_path, monkeypatch):
    """HVD_TIMELINE on the path users run (the graph-captured engine): the tail kernel stamps %globaltimer into a
    page-locked ring and the engine drains it lazily - a 10-step run leaves 10 FUSED_ALLREDUCE_UPDATE_LL and 10
    FUSED_ALLREDUCE_UPDATE_OWNER_ROWS complete events with plausible device durations."""
    import json
    _lib()
    path = tmp_path / "timeline.json"
    monkeypatch.setenv("HVD_TIMELINE", str(path))
    import horovod_distributed_example_b200 as _hvd
    _hvd.shutdown()                       # an earlier test may have left a world without a timeline behind
    hvd, build = _models(0.25, 0.5, "adadelta")
    m = build()
    rng = np.random.default_rng(2)
    x = rng.random((B, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, B)
    for _ in range(10):
        m.train_on_batch(x, y)
    assert type(m._backend).__name__ == "MnistEngine" and m._backend.pipe is not None
    assert hvd.basics.state().timeline is not None and m._backend._tl_active
    m._backend.drain_timeline(final=True)          # (shutdown() does this too; called here so an error surfaces)
    hvd.shutdown()
    ev = [e for e in json.loads(path.read_text()) if isinstance(e, dict)]
    ll = [e for e in ev if e.get("name") == "FUSED_ALLREDUCE_UPDATE_LL" and e.get("ph") == "X"]
    own = [e for e in ev if e.get("name") == "FUSED_ALLREDUCE_UPDATE_OWNER_ROWS" and e.get("ph") == "X"]
    assert len(ll) == 10 and len(own) == 10, (len(ll), len(own))
    assert all(0 < e["dur"] < 1000 for e in ll + own)                  # microseconds of device time
    ts = [e["ts"] for e in ll]
    assert ts == sorted(ts)


def test_engine_trains_with_dropout_and_partial_batch(monkeypatch):
    _lib()
    monkeypatch.setenv("HVD_DEBUG_GUARD", "1")      # red zones around every activation buffer
    hvd, build = _models(0.25, 0.5, "adadelta")
    from horovod_distributed_example_b200 import keras
    (x, y), (xt, yt) = keras.datasets.synthetic_mnist(1024 + 96, 256)
    x = (x[..., None] / 255.0).astype(np.float32)
    xt = (xt[..., None] / 255.0).astype(np.float32)
    m = build()
    h = m.fit(x, y, batch_size=128, epochs=6, verbose=0, validation_data=(xt, yt))
    assert type(m._backend).__name__ == "MnistEngine"
    assert h.history["loss"][-1] < 0.3 < h.history["loss"][0]            # the reference's CI range
    assert h.history["val_accuracy"][-1] > 0.9
    assert m.predict(xt[:37]).shape == (37, 10)
    assert m._backend.check_guards()                 # no kernel wrote outside its buffer (full and partial batches)
    hvd.shutdown()


def _dense_lib():
    lib = _lib()
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.hvd_dense_forward.restype = ci
    lib.hvd_dense_forward.argtypes = [vp, vp, vp, vp, ci, ci, ci, ci, ci, vp, vp]
    return lib


@pytest.mark.parametrize("M,N,K,relu,ks", [(128, 128, 9216, 1, 36), (100, 16, 128, 0, 1), (128, 64, 2048, 1, 8)])
def test_dense_forward_splitk_with_caller_scratch(M, N, K, relu, ks):
    lib = _dense_lib()
    torch.manual_seed(8)
    x, W, b = torch.randn(M, K, device="cuda"), torch.randn(K, N, device="cuda") * 0.05, torch.randn(N, device="cuda")
    y = torch.empty(M, N, device="cuda")
    scratch = torch.zeros(M * N, device="cuda")
    for _ in range(2):                                   # the finalize pass re-arms the scratch
        assert lib.hvd_dense_forward(x.data_ptr(), W.data_ptr(), b.data_ptr(), y.data_ptr(), M, N, K, relu, ks,
                                     scratch.data_ptr(), _s()) == 0
    ref = x.double() @ W.double() + b.double()
    ref = F.relu(ref) if relu else ref
    assert _rel(y, ref) < 3e-3
    assert (scratch == 0).all()


def test_async_host_steps_match_blocking_steps():
    """train_on_batch_async (pipelined H2D / D2H, inputs double-buffered) gives the same losses and
    weights as blocking train_on_batch on the same batches (dropout off, same seed)."""
    _lib()
    hvd, build = _models(0.0, 0.0, "adadelta")
    rng = np.random.default_rng(1)
    X = rng.random((6, B, 28, 28, 1), dtype=np.float32)
    Y = rng.integers(0, 10, (6, B))
    a, b = build(), build()
    la = [a.train_on_batch(X[i], Y[i])[0] for i in range(6)]
    hs = [b.train_on_batch_async(X[i], Y[i]) for i in range(6)]       # all six enqueued before any is read
    lb = [h.result()[0] for h in hs]
    np.testing.assert_allclose(la, lb, rtol=2e-3, atol=1e-4)          # red.add ordering differs run to run
    for wa, wb in zip(a.get_weights(), b.get_weights()):
        np.testing.assert_allclose(wa, wb, rtol=0, atol=5e-3)
    hvd.shutdown()


@pytest.mark.gpu
def test_fit_per_batch_logs_arrive_without_a_sync(tmp_path):
    """fit() on the fused engine hands callbacks LazyScalars: TensorBoard(update_freq='batch') still records every
    batch exactly once, in order, and the epoch loss equals the sample-weighted mean of those batch losses."""
    _lib()
    hvd, build = _models(0.25, 0.5, "adadelta")
    from horovod_distributed_example_b200 import keras
    from horovod_distributed_example_b200.utils import tfevents
    (x, y), _ = keras.datasets.synthetic_mnist(2048 + 40, 16)
    x = (x[..., None] / 255.0).astype(np.float32)
    m = build()
    seen = []
    cbs = [keras.callbacks.TensorBoard(log_dir=str(tmp_path), update_freq="batch"),
           keras.callbacks.LambdaCallback(on_batch_end=lambda b, logs: seen.append(type(logs["loss"]).__name__))]
    h = m.fit(x, y, batch_size=128, epochs=2, verbose=0, callbacks=cbs)
    assert type(m._backend).__name__ == "MnistEngine" and set(seen) == {"LazyScalar"}
    ev = tfevents.read_scalars(str(next((tmp_path / "train").iterdir())))
    bl = [(s, v) for t, s, v in ev if t == "batch_loss"]
    assert [s for s, _ in bl] == list(range(1, 35))                 # 17 steps x 2 epochs, each once, in order
    sizes = [128] * 16 + [40]
    for ep in range(2):
        vals = [v for _, v in bl[17 * ep:17 * (ep + 1)]]
        assert np.isfinite(vals).all()
        want = sum(v * n for v, n in zip(vals, sizes)) / sum(sizes)
        assert abs(h.history["loss"][ep] - want) < 1e-5 * max(1.0, abs(want))
    assert h.history["loss"][1] < h.history["loss"][0]
    hvd.shutdown()
