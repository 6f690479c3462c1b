"""MNIST CNN trained data-parallel with Adam, LR warmup and metric averaging - the capability of the
reference's `tensorflow2_keras_mnist.py` (per-rank dataset cache, repeat/shuffle/batch input pipeline,
Adam with LR 0.001 x ranks, Broadcast / MetricAverage / LearningRateWarmup callbacks in that order,
rank-0 checkpoints and TensorBoard logs, 500//ranks steps x 24 epochs) expressed with this framework.

`HVD_EXAMPLE_STEPS` / `HVD_EXAMPLE_EPOCHS` shrink the run for smoke tests.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras


def input_pipeline(rank: int) -> keras.Dataset:
    # one cache file per rank so concurrent ranks never race on the same download/cache path
    (images, labels), _ = keras.datasets.mnist.load_data(path=f"mnist-{rank}.npz")
    x = (images[..., np.newaxis] / 255.0).astype(np.float32)
    y = labels.astype(np.int64)
    return keras.Dataset.from_tensor_slices((x, y)).repeat().shuffle(10000).batch(128)


def build_model() -> keras.Sequential:
    L = keras.layers
    return keras.Sequential([
        L.Conv2D(32, [3, 3], activation="relu"), L.Conv2D(64, [3, 3], activation="relu"),
        L.MaxPooling2D(pool_size=(2, 2)), L.Dropout(0.25), L.Flatten(),
        L.Dense(128, activation="relu"), L.Dropout(0.5), L.Dense(10, activation="softmax")])


def main():
    model_dir, _export_dir = hvd.model_paths()
    hvd.init()                                   # pins the GPU for this local rank when there is one
    world, me = hvd.size(), hvd.rank()

    model = build_model()
    optimizer = hvd.DistributedOptimizer(keras.optimizers.Adam(0.001 * world))
    # (Keras' experimental_run_tf_function=False is accepted and ignored: the wrapper is always on the gradient path)
    model.compile(loss=keras.losses.SparseCategoricalCrossentropy(), optimizer=optimizer, metrics=["accuracy"],
                  experimental_run_tf_function=False)

    callbacks = [
        hvd.callbacks.BroadcastGlobalVariablesCallback(0),                     # identical start on every rank
        hvd.callbacks.MetricAverageCallback(),                                 # must precede metric consumers
        hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=3, verbose=1),  # lr/ranks -> lr over 3 epochs
    ]
    if me == 0:
        callbacks += [keras.callbacks.ModelCheckpoint(os.path.join(model_dir, "checkpoint-{epoch}.h5")),
                      keras.callbacks.TensorBoard(log_dir=model_dir, update_freq="batch")]

    steps = int(os.environ.get("HVD_EXAMPLE_STEPS", 500 // world))             # fewer steps with more GPUs
    epochs = int(os.environ.get("HVD_EXAMPLE_EPOCHS", 24))
    model.fit(input_pipeline(me), steps_per_epoch=steps, epochs=epochs, callbacks=callbacks, verbose=int(me == 0))


if __name__ == "__main__":
    main()
