import json
import struct

import numpy as np
import pytest

from horovod_distributed_example_b200 import keras
from horovod_distributed_example_b200.keras import optimizers as O
from horovod_distributed_example_b200.utils import hdf5


def test_roundtrip_groups_datasets_attrs(tmp_path):
    p = str(tmp_path / "t.h5")
    a = np.arange(24, dtype=np.float32).reshape(2, 3, 4)
    b = np.array(7, dtype=np.int64)
    with hdf5.Writer(p) as w:
        w.create_group("/g1/sub")
        w.create_dataset("/g1/sub/a:0", a)
        w.create_dataset("/scalar", b)
        w.create_dataset("/g1/empty", np.zeros((0, 3), np.float32))
        w.set_attr("/", "title", b"hello")
        w.set_attr("/g1", "names", [b"x", b"longer_name", b""])
        w.set_attr("/g1/sub/a:0", "scale", np.array([1.5, 2.5], np.float64))
        w.set_attr("/g1", "empty_list", [])
    with hdf5.Reader(p) as r:
        assert r.list("/") == ["g1", "scalar"]
        assert r.list("/g1") == ["empty", "sub"]
        assert r.kind("/g1/sub") == "group" and r.kind("/g1/sub/a:0") == "dataset" and r.kind("/nope") == "missing"
        np.testing.assert_array_equal(r.dataset("/g1/sub/a:0"), a)
        assert r.dataset("/scalar") == 7 and r.dataset("/scalar").dtype == np.int64
        assert r.dataset("/g1/empty").shape == (0, 3)
        assert r.attr("/", "title") == b"hello"
        assert r.attr("/g1", "names") == [b"x", b"longer_name", b""]
        np.testing.assert_array_equal(r.attr("/g1/sub/a:0", "scale"), [1.5, 2.5])


def test_file_structure_is_hdf5_v0(tmp_path):
    """Byte-level checks of the superblock and the root symbol-table entry."""
    p = str(tmp_path / "s.h5")
    with hdf5.Writer(p) as w:
        w.create_dataset("/d", np.ones(3, np.float32))
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n"
    assert raw[8] == 0 and raw[13] == 8 and raw[14] == 8                 # v0, 8-byte offsets/lengths
    eof = struct.unpack_from("<Q", raw, 40)[0]
    assert eof == len(raw)
    root_hdr, cache_type = struct.unpack_from("<Q", raw, 64)[0], struct.unpack_from("<I", raw, 72)[0]
    assert cache_type == 1 and raw[root_hdr] == 1                        # v1 object header
    btree, heap = struct.unpack_from("<QQ", raw, 80)
    assert raw[btree:btree + 4] == b"TREE" and raw[heap:heap + 4] == b"HEAP"
    assert struct.unpack_from("<Q", raw, heap + 16)[0] == 1              # H5HL_FREE_NULL
    snod = struct.unpack_from("<Q", raw, btree + 32)[0]
    assert raw[snod:snod + 4] == b"SNOD"
    assert len(raw) % 8 == 0


def test_many_entries_span_symbol_nodes(tmp_path):
    p = str(tmp_path / "m.h5")
    names = [f"w{i:04d}" for i in range(300)]
    with hdf5.Writer(p) as w:
        for i, n in enumerate(names):
            w.create_dataset("/g/" + n, np.full(2, i, np.float32))
    with hdf5.Reader(p) as r:
        assert r.list("/g") == sorted(names)
        assert r.dataset("/g/w0299")[0] == 299


def test_oversized_attribute_is_an_error(tmp_path):
    w = hdf5.Writer(str(tmp_path / "big.h5"))
    w.set_attr("/", "blob", b"x" * 70000)
    with pytest.raises(hdf5.H5Error):
        w.close()


def _mnist(seed=0):
    keras.backend.clear_session()
    return keras.Sequential([
        keras.layers.Conv2D(32, (3, 3), activation='relu', input_shape=(28, 28, 1)),
        keras.layers.Conv2D(64, (3, 3), activation='relu'),
        keras.layers.MaxPooling2D((2, 2)), keras.layers.Dropout(0.25), keras.layers.Flatten(),
        keras.layers.Dense(128, activation='relu'), keras.layers.Dropout(0.5),
        keras.layers.Dense(10, activation='softmax')], seed=seed)


def test_keras_checkpoint_layout(tmp_path, hvd_single):
    m = _mnist()
    m.compile(optimizer=hvd_single.DistributedOptimizer(O.Adadelta(1.0)),
              loss="categorical_crossentropy", metrics=["accuracy"])
    x = np.random.default_rng(0).random((8, 28, 28, 1), dtype=np.float32)
    y = keras.utils.to_categorical(np.arange(8) % 10, 10)
    m.train_on_batch(x, y)
    p = str(tmp_path / "checkpoint-1.h5")
    m.save(p)
    with hdf5.Reader(p) as r:
        assert set(r.attr_names("/")) >= {"keras_version", "backend", "model_config", "training_config"}
        assert r.attr("/", "backend") == b"tensorflow"
        cfg = json.loads(r.attr("/", "model_config"))
        assert cfg["class_name"] == "Sequential"
        assert [l["class_name"] for l in cfg["config"]["layers"]] == [
            "Conv2D", "Conv2D", "MaxPooling2D", "Dropout", "Flatten", "Dense", "Dropout", "Dense"]
        tc = json.loads(r.attr("/", "training_config"))
        assert tc["optimizer_config"]["class_name"] == "Adadelta" and tc["loss"] == "categorical_crossentropy"
        names = [b.decode() for b in r.attr("/model_weights", "layer_names")]
        assert names == ["conv2d_1", "conv2d_2", "max_pooling2d_1", "dropout_1", "flatten_1",
                         "dense_1", "dropout_2", "dense_2"]
        assert r.attr("/model_weights/conv2d_1", "weight_names") == [b"conv2d_1/kernel:0", b"conv2d_1/bias:0"]
        assert r.attr("/model_weights/flatten_1", "weight_names") == []
        shapes = {n: r.dataset(f"/model_weights/{l}/{l}/{n}:0").shape for l, n in
                  [("conv2d_1", "kernel"), ("conv2d_2", "kernel"), ("dense_1", "kernel"), ("dense_2", "kernel"),
                   ("conv2d_1", "bias"), ("dense_2", "bias")]}
        assert r.dataset("/model_weights/conv2d_1/conv2d_1/kernel:0").shape == (3, 3, 1, 32)      # HWIO
        assert r.dataset("/model_weights/conv2d_2/conv2d_2/kernel:0").shape == (3, 3, 32, 64)
        assert r.dataset("/model_weights/dense_1/dense_1/kernel:0").shape == (9216, 128)
        assert r.dataset("/model_weights/dense_2/dense_2/bias:0").shape == (10,)
        ow = r.attr("/optimizer_weights", "weight_names")
        assert len(ow) == 16                                                                 # 8 acc + 8 delta acc
        assert r.dataset("/optimizer_weights/" + ow[0].decode()).dtype == np.float32
    assert m.count_params() == 1199882


def test_save_load_resume_is_exact(tmp_path, hvd_single):
    m = _mnist(seed=3)
    m.compile(optimizer=O.Adam(0.01), loss=keras.losses.SparseCategoricalCrossentropy(), metrics=["accuracy"])
    rng = np.random.default_rng(0)
    x = rng.random((16, 28, 28, 1), dtype=np.float32); y = rng.integers(0, 10, 16)
    for _ in range(3):
        m.train_on_batch(x, y)
    p = str(tmp_path / "m.h5")
    m.save(p)
    m2 = keras.load_model(p)
    assert type(m2.optimizer).__name__ == "Adam" and m2.optimizer.iterations == 3
    np.testing.assert_array_equal(m.predict(x), m2.predict(x))
    for a, b in zip(m.arena.slots, m2.arena.slots):
        np.testing.assert_array_equal(a.numpy(), b.numpy())
    # weights-only file + load_weights
    pw = str(tmp_path / "w.h5")
    m.save_weights(pw)
    m3 = _mnist(seed=9); m3.build()
    m3.load_weights(pw)
    np.testing.assert_array_equal(m.predict(x), m3.predict(x))
