"""BASELINE.json config 1: the mnist_keras.py CNN, single process on CPU, batch 128, synthetic data."""
import os

import numpy as np
import torch

from horovod_distributed_example_b200 import keras
from horovod_distributed_example_b200.keras import optimizers as O
from horovod_distributed_example_b200.utils import metrics_sink, tfevents


def _data(n=1024):
    (x, y), (xt, yt) = keras.datasets.synthetic_mnist(n, 256)
    return (x[..., None] / 255.0).astype(np.float32), y, (xt[..., None] / 255.0).astype(np.float32), yt


def _model(seed=0):
    keras.backend.clear_session()
    m = keras.Sequential(seed=seed)
    m.add(keras.layers.Conv2D(32, kernel_size=(3, 3), activation='relu', input_shape=(28, 28, 1)))
    m.add(keras.layers.Conv2D(64, (3, 3), activation='relu'))
    m.add(keras.layers.MaxPooling2D(pool_size=(2, 2)))
    m.add(keras.layers.Dropout(0.25))
    m.add(keras.layers.Flatten())
    m.add(keras.layers.Dense(128, activation='relu'))
    m.add(keras.layers.Dropout(0.5))
    m.add(keras.layers.Dense(10, activation='softmax'))
    return m


def test_shapes_and_param_count():
    m = _model(); m.build()
    assert [l.output_shape for l in m.layers] == [(26, 26, 32), (24, 24, 64), (12, 12, 64), (12, 12, 64),
                                                  (9216,), (128,), (128,), (10,)]
    assert m.count_params() == 320 + 18496 + 1179776 + 1290 == 1199882
    assert m.arena.n_params() * 4 == 4799528
    # arena is in backward-ready order: dense_2 first, conv2d_1 last
    assert [s.name for s in m.arena.specs][:2] == ["dense_2/bias", "dense_2/kernel"]
    assert m.arena.specs[-1].name == "conv2d_1/kernel"


def test_flatten_order_is_hwc():
    m = _model(); m.build()
    x = torch.zeros(1, 12, 12, 64); x[0, 2, 3, 5] = 1.0
    assert int(m.layers[4].call(x, False).argmax()) == 2 * (12 * 64) + 3 * 64 + 5


def test_tf1_script_flow_reaches_ci_threshold(tmp_path, hvd_single):
    x, y, xt, yt = _data(1536)
    yc, ytc = keras.utils.to_categorical(y, 10), keras.utils.to_categorical(yt, 10)
    metrics_sink.init(True, str(tmp_path / "metrics.jsonl"))
    m = _model(seed=1)
    opt = hvd_single.DistributedOptimizer(O.Adadelta(1.0 * hvd_single.size()))
    m.compile(loss=keras.losses.categorical_crossentropy, optimizer=opt, metrics=['accuracy'])
    cbs = [hvd_single.callbacks.BroadcastGlobalVariablesCallback(0),
           keras.callbacks.ModelCheckpoint(str(tmp_path / "checkpoint-{epoch}.h5")),
           keras.callbacks.TensorBoard(log_dir=str(tmp_path / "eval"), update_freq='batch')]
    h = m.fit(x, yc, batch_size=128, epochs=3, verbose=0, callbacks=cbs, validation_data=(xt, ytc))
    score = m.evaluate(xt, ytc, verbose=0)
    assert h.history["loss"][-1] < 0.3 < h.history["loss"][0]          # the reference's CI range
    assert score[0] < 0.3 and score[1] > 0.9
    assert set(h.history) == {"loss", "accuracy", "val_loss", "val_accuracy"}
    assert sorted(f for f in os.listdir(tmp_path) if f.endswith(".h5")) == [
        "checkpoint-1.h5", "checkpoint-2.h5", "checkpoint-3.h5"]
    ev = [f for f in os.listdir(tmp_path / "eval" / "train")]
    sc = tfevents.read_scalars(str(tmp_path / "eval" / "train" / ev[0]))
    assert sum(t == "batch_loss" for t, _, _ in sc) == 3 * 12 and any(t == "epoch_loss" for t, _, _ in sc)
    assert metrics_sink.check("epoch_loss", "0.0..0.3", "last")["passed"]
    metrics_sink.init(False)


def test_partial_last_batch_and_eval_weighting(hvd_single):
    x, y, _, _ = _data(300)
    m = _model(seed=2)
    m.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=O.Adam(1e-3), metrics=['accuracy'])
    seen = []
    m.fit(x, y, batch_size=128, epochs=1, verbose=0, callbacks=[
        keras.callbacks.LambdaCallback(on_batch_end=lambda b, logs: seen.append(logs["size"]))])
    assert seen == [128, 128, 44]
    full = m.evaluate(x, y, batch_size=300, verbose=0)
    parts = m.evaluate(x, y, batch_size=128, verbose=0)
    np.testing.assert_allclose(full, parts, rtol=1e-5)


def test_losses_match_keras_definition():
    p = torch.tensor([[0.7, 0.2, 0.1], [0.0, 1.0, 0.0]])
    t = torch.tensor([[1.0, 0, 0], [1.0, 0, 0]])
    l = keras.losses.categorical_crossentropy(t, p)
    np.testing.assert_allclose(l.numpy(), [-np.log(0.7), -np.log(1e-7)], rtol=1e-5)
    ls = keras.losses.sparse_categorical_crossentropy(torch.tensor([0, 0]), p)
    np.testing.assert_allclose(ls.numpy(), l.numpy(), rtol=1e-6)


def test_export_serving_artifacts(tmp_path, hvd_single):
    from horovod_distributed_example_b200.utils import saved_model
    x, y, _, _ = _data(64)
    m = _model(seed=4)
    m.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=O.Adadelta(1.0), metrics=['accuracy'])
    m.train_on_batch(x, y)
    out = saved_model.export(m, str(tmp_path / "20260101-000000"), inputs={'input': None}, outputs={'prob': None})
    assert sorted(os.listdir(out)) == ["model.h5", "model.ts", "signature.json"]
    ts = torch.jit.load(os.path.join(out, "model.ts"))
    np.testing.assert_allclose(ts(torch.from_numpy(x)).numpy(), m.predict(x), rtol=1e-5, atol=1e-6)


def test_torchmodel_resnet_tiny_trains_through_arena(hvd_single):
    """Any nn.Module (here a tiny ResNet with BatchNorm and residuals) trains through the same arena /
    DistributedOptimizer path: parameters are re-homed into the arena in reverse registration order."""
    from horovod_distributed_example_b200.models.resnet import resnet_tiny
    torch.manual_seed(0)
    net = resnet_tiny(10)
    n_params = sum(p.numel() for p in net.parameters())
    m = keras.TorchModel(net)
    m.compile(optimizer=hvd_single.DistributedOptimizer(O.SGD(0.05, momentum=0.9)),
              loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    rng = np.random.default_rng(0)
    x = rng.random((32, 32, 32, 3), dtype=np.float32)
    y = (x.mean(axis=(1, 2, 3)) > 0.5).astype(np.int64)
    losses = [m.train_on_batch(x, y)[0] for _ in range(12)]
    assert m.arena.n_params() == n_params and losses[-1] < losses[0]
    assert m.arena.specs[0].name.startswith("fc.")                  # last layer first: backward-ready order
    for name, p in net.named_parameters():
        assert p.data_ptr() == m.arena.param(name).data_ptr()       # module weights ARE arena views
    from horovod_distributed_example_b200.parallel import buckets as B
    bk = B.plan(m.arena, 16 << 10)
    assert len(bk) > 3 and bk[0].start == 0 and bk[-1].end == m.arena.n


def test_channels_first_matches_channels_last():
    """The reference's other input branch (mnist_keras.py:50-53): the same weights give the same outputs whether
    the images arrive as (N,1,28,28) or (N,28,28,1); Flatten keeps the (h,w,c) order in both."""
    rng = np.random.default_rng(3)
    x = rng.random((4, 28, 28, 1), dtype=np.float32)
    m_last = _model(seed=5); m_last.build()
    keras.backend.set_image_data_format("channels_first")
    try:
        assert keras.backend.image_data_format() == "channels_first"
        keras.backend.clear_session()
        m_first = keras.Sequential([
            keras.layers.Conv2D(32, (3, 3), activation='relu', input_shape=(1, 28, 28)),
            keras.layers.Conv2D(64, (3, 3), activation='relu'), keras.layers.MaxPooling2D((2, 2)),
            keras.layers.Dropout(0.25), keras.layers.Flatten(), keras.layers.Dense(128, activation='relu'),
            keras.layers.Dropout(0.5), keras.layers.Dense(10, activation='softmax')], seed=5)
        m_first.build()
        assert [l.output_shape for l in m_first.layers][:3] == [(32, 26, 26), (64, 24, 24), (64, 12, 12)]
        m_first.set_weights(m_last.get_weights())
        p1 = m_first.predict(np.ascontiguousarray(x.transpose(0, 3, 1, 2)))
    finally:
        keras.backend.set_image_data_format("channels_last")
    np.testing.assert_allclose(p1, m_last.predict(x), rtol=1e-5, atol=1e-6)
    assert m_first.get_config()["layers"][0]["config"]["data_format"] == "channels_first"
