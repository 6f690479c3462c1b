"""Worker bodies for the CPU/gloo world_size=2 tests (BASELINE.json config 2)."""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras


def _mnist_model(seed):
    return keras.Sequential([
        keras.layers.Conv2D(32, [3, 3], activation='relu'),
        keras.layers.Conv2D(64, [3, 3], activation='relu'),
        keras.layers.MaxPooling2D(pool_size=(2, 2)),
        keras.layers.Dropout(0.25),
        keras.layers.Flatten(),
        keras.layers.Dense(128, activation='relu'),
        keras.layers.Dropout(0.5),
        keras.layers.Dense(10, activation='softmax')], seed=seed)


def collectives():
    hvd.init()
    r, n = hvd.rank(), hvd.size()
    t = torch.full((5,), float(r + 1))
    assert torch.allclose(hvd.allreduce(t, average=True), torch.full((5,), (n + 1) / 2))
    assert torch.allclose(hvd.allreduce(t, average=False), torch.full((5,), n * (n + 1) / 2))
    b = hvd.broadcast(torch.arange(4.0) * (r + 1), root_rank=1)
    assert torch.equal(b, torch.arange(4.0) * 2)
    g = hvd.allgather(torch.full((r + 1, 2), float(r)))
    assert g.shape == (sum(range(1, n + 1)), 2)
    assert hvd.broadcast_object({"a": r}, 0) == {"a": 0}
    assert hvd.allreduce(float(r)) == (n - 1) / 2
    print("OK")


def broadcast_and_grads():
    """Rank 0's weights arrive everywhere; averaged gradients equal the mean of per-rank
    gradients; replicas stay identical after a step."""
    hvd.init()
    r, n = hvd.rank(), hvd.size()
    m = _mnist_model(seed=100 + r)           # different init per rank on purpose
    opt = hvd.DistributedOptimizer(keras.optimizers.Adadelta(1.0 * n))
    assert type(opt).__name__ == "Adadelta"
    m.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=opt, metrics=['accuracy'])
    rng = np.random.default_rng(7 + r)       # different data per rank
    x = rng.random((16, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, 16)
    m._ensure_built(x.shape)
    w_before = hvd.allgather(m.arena.params.clone()[None])
    assert not torch.equal(w_before[0], w_before[1])
    cb = hvd.callbacks.BroadcastGlobalVariablesCallback(0)
    cb.set_model(m)
    cb.on_train_begin()
    w_after = hvd.allgather(m.arena.params.clone()[None])
    assert torch.equal(w_after[0], w_after[1]) and torch.equal(w_after[0], w_before[0])
    # local grads, then compare the averaged grads with the mean computed from a gather
    be = m._backend
    m.arena.grads.zero_()
    with torch.enable_grad():
        yp = m._forward(torch.from_numpy(x), False)
        loss = m.loss(torch.from_numpy(y), yp).mean()
    loss.backward()
    local = m.arena.grads.clone()
    gathered = hvd.allgather(local[None])
    avg = opt.get_gradients(m.arena).clone()
    assert torch.allclose(avg, gathered.mean(0), rtol=1e-6, atol=1e-8)
    # one real step keeps replicas bit-identical
    m.train_on_batch(x, y)
    w2 = hvd.allgather(m.arena.params.clone()[None])
    assert torch.equal(w2[0], w2[1])
    print("OK")


def tf2_script_flow():
    """The TF2 script's callback stack at ws=2: metric average leaves identical logs on
    both ranks, lr is not averaged, only rank 0 writes files."""
    hvd.init()
    r, n = hvd.rank(), hvd.size()
    out = os.environ["HVD_TEST_OUT"]
    (xi, yi), _ = keras.datasets.mnist.load_data(path=os.path.join(out, f"mnist-{r}.npz"), n_train=1024, n_test=64)
    ds = keras.Dataset.from_tensor_slices(((xi[..., None] / 255.0).astype(np.float32), yi.astype(np.int64)))
    ds = ds.repeat().shuffle(512, seed=r).batch(32)
    m = _mnist_model(seed=r)
    opt = hvd.DistributedOptimizer(keras.optimizers.Adam(0.001 * n))
    m.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=opt, metrics=['accuracy'],
              experimental_run_tf_function=False)
    seen = []
    cbs = [hvd.callbacks.BroadcastGlobalVariablesCallback(0),
           hvd.callbacks.MetricAverageCallback(),
           hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=2, verbose=0),
           keras.callbacks.LambdaCallback(on_epoch_end=lambda e, logs: seen.append(dict(logs)))]
    if r == 0:
        cbs.append(keras.callbacks.ModelCheckpoint(os.path.join(out, "ckpt", "checkpoint-{epoch}.h5")))
        cbs.append(keras.callbacks.TensorBoard(log_dir=os.path.join(out, "tb"), update_freq='batch'))
    m.fit(ds, steps_per_epoch=4, epochs=3, callbacks=cbs, verbose=0)
    all_seen = hvd.allgather_object(seen)
    assert all_seen[0] == all_seen[1], (all_seen[0], all_seen[1])
    assert abs(seen[-1]["lr"] - 0.001 * n) < 1e-9
    assert seen[0]["lr"] < seen[1]["lr"] <= seen[2]["lr"]
    w = hvd.allgather(m.arena.params.clone()[None])
    assert torch.equal(w[0], w[1])
    hvd.barrier()
    if r == 0:
        assert sorted(os.listdir(os.path.join(out, "ckpt"))) == ["checkpoint-1.h5", "checkpoint-2.h5", "checkpoint-3.h5"]
    print("OK")


def resume_from_checkpoint():
    """Resume path: rank 0 restores a checkpoint, the broadcast callback makes
    every rank identical to it (weights, optimizer slots, iteration counter), training continues."""
    hvd.init()
    r = hvd.rank()
    path = os.path.join(os.environ["HVD_TEST_OUT"], "ckpt.h5")
    rng = np.random.default_rng(5)
    x = rng.random((64, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, 64)
    if r == 0:
        m0 = _mnist_model(seed=1)
        m0.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=keras.optimizers.Adam(0.01))
        for _ in range(3):
            m0.train_on_batch(x, y)
        m0.save(path)
    hvd.barrier()
    m = _mnist_model(seed=50 + r)
    opt = hvd.DistributedOptimizer(keras.optimizers.Adam(0.01))
    m.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=opt, metrics=['accuracy'])
    m._ensure_built(x.shape)
    if r == 0:
        restored = keras.load_model(path)                 # rank-0-only load, as in the reference
        m.set_weights(restored.get_weights())
        for a, b in zip(m.arena.slots, restored.arena.slots):
            a.copy_(b)
        opt.iterations = restored.optimizer.iterations
    cb = hvd.callbacks.BroadcastGlobalVariablesCallback(0)
    cb.set_model(m)
    cb.on_train_begin()
    st = hvd.allgather(torch.cat([m.arena.params, m.arena.slots[0], m.arena.slots[1]])[None])
    assert torch.equal(st[0], st[1])
    assert opt.iterations == 3
    m.train_on_batch(x, y)
    assert opt.iterations == 4
    w = hvd.allgather(m.arena.params.clone()[None])
    assert torch.equal(w[0], w[1])
    print("OK")


def hvd_load_model_and_optimizer_state():
    """hvd.load_model on every rank (Horovod's resume recipe): the optimizer comes back wrapped, with its state;
    broadcast_optimizer_state equalises keras-style and torch.optim optimizers; timeline control at run time."""
    hvd.init()
    r = hvd.rank()
    out = os.environ["HVD_TEST_OUT"]
    path = os.path.join(out, "ckpt.h5")
    rng = np.random.default_rng(5)
    x = rng.random((32, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, 32)
    if r == 0:
        m0 = _mnist_model(seed=1)
        m0.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=keras.optimizers.Adam(0.005))
        for _ in range(2):
            m0.train_on_batch(x, y)
        m0.save(path)
        ref_slots = [s.clone() for s in m0.arena.slots]
    hvd.barrier()
    hvd.start_timeline(os.path.join(out, "tl.json"))
    m = hvd.load_model(path)                              # every rank
    opt = m.optimizer
    assert type(opt).__name__ == "Adam" and getattr(opt, "_hvd_distributed", False)
    assert opt.iterations == 2 and abs(opt.lr - 0.005) < 1e-9      # Adam stores its iteration counter in the file
    if r == 0:
        for a, b in zip(m.arena.slots, ref_slots):
            assert torch.equal(a, b)
    # perturb rank 1's state, then equalise
    if r == 1:
        for s in m.arena.slots:
            s.add_(1.0)
        opt.iterations = 9
    hvd.broadcast_optimizer_state(opt, 0, model=m)
    st = hvd.allgather(torch.cat(list(m.arena.slots))[None])
    assert torch.equal(st[0], st[1]) and opt.iterations == 2
    m.train_on_batch(x, y)                                # continues training with averaged gradients
    w = hvd.allgather(m.arena.params.clone()[None])
    assert torch.equal(w[0], w[1])
    # torch.optim flavour
    lin = torch.nn.Linear(4, 3)
    topt = torch.optim.Adam(lin.parameters(), lr=0.01 * (r + 1))
    lin(torch.randn(5, 4) + r).sum().backward()
    topt.step()
    hvd.broadcast_parameters(lin.state_dict(), 0)
    hvd.broadcast_optimizer_state(topt, 0)
    ea = hvd.allgather(topt.state_dict()["state"][0]["exp_avg"].reshape(1, -1))
    assert torch.equal(ea[0], ea[1]) and abs(topt.param_groups[0]["lr"] - 0.01) < 1e-12
    hvd.stop_timeline()
    if r == 0:
        assert json.load(open(os.path.join(out, "tl.json")))   # a valid Chrome trace with the broadcasts in it
    print("OK")


def plan_mismatch():
    hvd.init()
    layers = [keras.layers.Flatten(), keras.layers.Dense(4 + hvd.rank())]
    m = keras.Sequential(layers, seed=0)
    m.compile(loss="mse", optimizer=hvd.DistributedOptimizer(keras.optimizers.SGD(0.1)))
    try:
        m._ensure_built((2, 3, 3, 1))
    except RuntimeError as e:
        assert "bucket plan mismatch" in str(e)
        print("OK")
        return
    raise AssertionError("mismatch not detected")


if __name__ == "__main__":
    globals()[sys.argv[1]]()
