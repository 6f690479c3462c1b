import math

import numpy as np
import pytest
import torch

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200.keras import optimizers as O
from horovod_distributed_example_b200.parallel.arena import ParamArena


def _arena(opt, n=300, seed=0):
    a = ParamArena([("w", (n,))], "cpu", n_slots=opt.n_slots)
    torch.manual_seed(seed)
    a.param("w").copy_(torch.randn(n))
    opt.bind(a)
    return a


def test_adadelta_matches_keras_formula():
    opt = O.Adadelta(lr=2.0)
    a = _arena(opt)
    p = a.param("w").numpy().astype(np.float64).copy()
    acc = np.zeros_like(p); dacc = np.zeros_like(p)
    rng = np.random.default_rng(1)
    for _ in range(5):
        g = rng.normal(size=p.shape).astype(np.float32)
        a.grad("w").copy_(torch.from_numpy(g))
        opt.apply_gradients(a)
        g64 = g.astype(np.float64)
        acc = 0.95 * acc + 0.05 * g64 ** 2
        u = g64 * np.sqrt(dacc + 1e-7) / np.sqrt(acc + 1e-7)
        p = p - 2.0 * u
        dacc = 0.95 * dacc + 0.05 * u ** 2
    np.testing.assert_allclose(a.param("w").numpy(), p, rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(a.slot(0, "w").numpy(), acc, rtol=2e-5)
    np.testing.assert_allclose(a.slot(1, "w").numpy(), dacc, rtol=2e-4, atol=1e-9)
    assert opt.iterations == 5


def test_adam_matches_tfkeras_formula():
    opt = O.Adam(lr=0.004)
    a = _arena(opt)
    p = a.param("w").numpy().astype(np.float64).copy()
    m = np.zeros_like(p); v = np.zeros_like(p)
    rng = np.random.default_rng(2)
    for t in range(1, 6):
        g = rng.normal(size=p.shape).astype(np.float32)
        a.grad("w").copy_(torch.from_numpy(g))
        opt.apply_gradients(a)
        g64 = g.astype(np.float64)
        lr_t = 0.004 * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        m = 0.9 * m + 0.1 * g64
        v = 0.999 * v + 0.001 * g64 ** 2
        p = p - lr_t * m / (np.sqrt(v) + 1e-7)
    np.testing.assert_allclose(a.param("w").numpy(), p, rtol=2e-5, atol=1e-6)


def test_lr_is_runtime_settable_and_reaches_device_vector():
    opt = O.Adadelta(1.0)
    _arena(opt)
    opt.lr = 0.25
    assert opt.lr == 0.25 and abs(float(opt.hyper[O.H_LR]) - 0.25) < 1e-7


def test_distributed_optimizer_keeps_class_name_and_config(hvd_single):
    base = O.Adadelta(3.0, rho=0.9)
    d = hvd_single.DistributedOptimizer(base)
    assert type(d).__name__ == "Adadelta" and isinstance(d, O.Adadelta)
    assert d.get_config() == base.get_config()
    assert O.serialize(d)["class_name"] == "Adadelta"
    with pytest.raises(ValueError):
        hvd_single.DistributedOptimizer(d)


def test_size1_skips_allreduce(hvd_single):
    d = hvd_single.DistributedOptimizer(O.SGD(0.5))
    a = _arena(d)
    before = a.param("w").clone()
    a.grad("w").fill_(1.0)
    d.apply_gradients(a)
    assert torch.allclose(a.param("w"), before - 0.5)


def test_backward_passes_per_step_accumulates(hvd_single):
    """DistributedOptimizer(backward_passes_per_step=2): two batches accumulate into ONE update that equals
    a single step on the sum of their gradients (round-1 bug: the slab was cleared on every pass)."""
    from horovod_distributed_example_b200 import keras

    def run(bpps, batches):
        keras.backend.clear_session()
        m = keras.Sequential([keras.layers.Dense(2, use_bias=False, input_shape=(2,))], seed=3)
        opt = hvd_single.DistributedOptimizer(O.SGD(lr=1.0), backward_passes_per_step=bpps)
        m.compile(loss=keras.losses.mean_squared_error, optimizer=opt)
        m.build()
        m.set_weights([np.zeros((2, 2), np.float32)])
        for x, y in batches:
            m.train_on_batch(x, y)
        return m.get_weights()[0]

    x1, y1 = np.array([[1., 0.]], np.float32), np.array([[1., 1.]], np.float32)
    x2, y2 = np.array([[0., 1.]], np.float32), np.array([[1., 1.]], np.float32)
    w_acc = run(2, [(x1, y1), (x2, y2)])
    # oracle: weights stay at zero during both passes, so the accumulated gradient is g(x1) + g(x2) at w = 0
    w1 = run(1, [(x1, y1)])
    w2 = run(1, [(x2, y2)])
    np.testing.assert_allclose(w_acc, w1 + w2, rtol=1e-6, atol=1e-7)
    assert np.abs(w_acc[0]).sum() > 0 and np.abs(w_acc[1]).sum() > 0      # neither pass was dropped
