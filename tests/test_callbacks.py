import pytest
import torch

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import basics, keras
from horovod_distributed_example_b200.keras import optimizers as O


class _FakeModel:
    def __init__(self, opt):
        self.optimizer = opt


@pytest.mark.parametrize("size", [1, 2, 8])
def test_warmup_closed_form(hvd_single, monkeypatch, size):
    monkeypatch.setattr(basics._state, "size", size)
    steps, warm = 10, 3
    opt = O.Adam(0.001 * size)
    cb = hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=warm, verbose=0)
    cb.set_model(_FakeModel(opt))
    cb.set_params({"steps": steps})
    cb.on_train_begin()
    init_lr = 0.001 * size
    lrs = []
    for epoch in range(5):
        cb.on_epoch_begin(epoch)
        for b in range(steps):
            cb.on_batch_begin(b)
            lrs.append(opt.lr)
            if epoch < warm:
                e = epoch + b / steps + 1.0 / steps
                expect = init_lr * (1.0 / size) * (e * (size - 1) / warm + 1)
                assert abs(opt.lr - expect) < 1e-12
            cb.on_batch_end(b)
        logs = {}
        cb.on_epoch_end(epoch, logs)
        assert logs["lr"] == opt.lr
    # ends exactly at the scaled LR and stays there; starts near lr/size
    assert abs(lrs[warm * steps - 1] - init_lr) < 1e-12 and abs(lrs[-1] - init_lr) < 1e-12
    assert abs(lrs[0] - init_lr / size * (1 + (size - 1) / (warm * steps))) < 1e-12
    monkeypatch.setattr(basics._state, "size", 1)


def test_momentum_correction_and_restore(hvd_single, monkeypatch):
    monkeypatch.setattr(basics._state, "size", 4)
    opt = O.SGD(0.4, momentum=0.9)
    cb = hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=2, steps_per_epoch=5)
    cb.set_model(_FakeModel(opt)); cb.set_params({"steps": 5})
    cb.on_train_begin(); cb.on_epoch_begin(0)
    old = opt.lr
    cb.on_batch_begin(0)
    assert abs(opt.momentum - 0.9 * opt.lr / old) < 1e-12
    cb.on_batch_end(0)
    assert opt.momentum == 0.9
    monkeypatch.setattr(basics._state, "size", 1)


def test_warmup_prints_finished_message(hvd_single, capsys):
    opt = O.Adam(0.001)
    cb = hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=1, verbose=1, steps_per_epoch=2)
    cb.set_model(_FakeModel(opt)); cb.set_params({"steps": 2})
    cb.on_train_begin(); cb.on_epoch_begin(0); cb.on_batch_begin(0); cb.on_batch_end(0)
    cb.on_epoch_end(0, {})
    assert "finished gradual learning rate warmup to 0.001" in capsys.readouterr().out


def test_metric_average_in_place_sorted(hvd_single, monkeypatch):
    """Emulate 2 ranks: the callback must visit metrics in sorted order, reduce them in one
    packed vector, and edit `logs` in place."""
    from horovod_distributed_example_b200 import callbacks as C
    monkeypatch.setattr(basics._state, "size", 2)
    calls = []

    def fake_allreduce_(vec, average=True, name=None, op=None):
        calls.append(vec.clone())
        vec.mul_(0.5).add_(0.5 * torch.tensor([10.0, 20.0, 30.0], dtype=vec.dtype))
        return vec
    monkeypatch.setattr(C.collectives, "allreduce_", fake_allreduce_)
    logs = {"val_loss": 3.0, "accuracy": 1.0, "loss": torch.tensor(2.0)}
    ref = logs
    C.MetricAverageCallback().on_epoch_end(0, logs)
    assert logs is ref and len(calls) == 1
    assert calls[0].tolist() == [1.0, 2.0, 3.0]            # sorted: accuracy, loss, val_loss
    assert logs == {"accuracy": 5.5, "loss": 11.0, "val_loss": 16.5}
    monkeypatch.setattr(basics._state, "size", 1)


def test_callback_order_lr_not_averaged(hvd_single):
    x = torch.rand(64, 4, 4, 1).numpy(); y = (torch.arange(64) % 3).numpy()
    m = keras.Sequential([keras.layers.Flatten(), keras.layers.Dense(3, activation="softmax")], seed=0)
    m.compile(optimizer=hvd.DistributedOptimizer(O.SGD(0.1)),
              loss=keras.losses.SparseCategoricalCrossentropy(), metrics=["accuracy"])
    seen = {}
    h = m.fit(x, y, batch_size=16, epochs=2, verbose=0, callbacks=[
        hvd.callbacks.BroadcastGlobalVariablesCallback(0), hvd.callbacks.MetricAverageCallback(),
        hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=1),
        keras.callbacks.LambdaCallback(on_epoch_end=lambda e, logs: seen.update(logs))])
    assert set(seen) == {"loss", "accuracy", "lr"} and h.history["lr"] == [0.1, 0.1]


def test_lr_scheduler_csv_logger_and_terminate_on_nan(tmp_path, hvd_single):
    import numpy as np
    from horovod_distributed_example_b200 import keras
    rng = np.random.default_rng(0)
    x = rng.normal(size=(96, 6)).astype(np.float32)
    y = (x[:, 0] > 0).astype(np.int64)
    m = keras.Sequential([keras.layers.Dense(8, activation="relu", input_shape=(6,)), keras.layers.Dense(2, activation="softmax")])
    m.compile(loss="sparse_categorical_crossentropy", optimizer=keras.optimizers.SGD(0.1), metrics=["accuracy"])
    csv = tmp_path / "log.csv"
    h = m.fit(x, y, batch_size=32, epochs=3, verbose=0,
              callbacks=[keras.callbacks.LearningRateScheduler(lambda e, lr: 0.1 * 0.5 ** e), keras.callbacks.CSVLogger(str(csv))])
    assert h.history["lr"] == [0.1, 0.05, 0.025] and abs(m.optimizer.lr - 0.025) < 1e-12
    rows = csv.read_text().strip().splitlines()
    assert rows[0] == "epoch,accuracy,loss,lr" and len(rows) == 4 and rows[3].startswith("2,")
    # a NaN loss stops training at the end of the epoch at the latest
    m2 = keras.Sequential([keras.layers.Dense(2, activation="softmax", input_shape=(6,))])
    m2.compile(loss="sparse_categorical_crossentropy", optimizer=keras.optimizers.SGD(0.1))
    xb = x.copy()
    xb[5, 2] = np.nan
    h2 = m2.fit(xb, y, batch_size=32, epochs=5, verbose=0, shuffle=False, callbacks=[keras.callbacks.TerminateOnNaN()])
    assert len(h2.history["loss"]) == 1 and m2.stop_training
