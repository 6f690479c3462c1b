"""MNIST CNN trained data-parallel with Adadelta - the capability of the reference's `mnist_keras.py`
(same model, batch size, epoch budget, LR scaling, callbacks, rank-0 checkpoints/logs, final
evaluate + save + reload + serving export) expressed with this framework.

    python examples/mnist_keras.py                                               # one process
    python -m horovod_distributed_example_b200.run -np 8 examples/mnist_keras.py  # one process per GPU

`HVD_EXAMPLE_SAMPLES` / `HVD_EXAMPLE_EPOCHS` shrink the run for smoke tests.
"""
import math
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras
from horovod_distributed_example_b200.utils import metrics_sink, saved_model

BATCH = 128
CLASSES = 10
BASE_EPOCHS = 12


def load_data():
    """MNIST as float32 NHWC in [0,1] with one-hot labels (synthetic MNIST-shaped data offline)."""
    (xtr, ytr), (xte, yte) = keras.datasets.mnist.load_data()
    limit = int(os.environ.get("HVD_EXAMPLE_SAMPLES", 0))
    if limit:
        xtr, ytr = xtr[:limit], ytr[:limit]
        xte, yte = xte[:max(limit // 4, 256)], yte[:max(limit // 4, 256)]

    def prep(x):
        # both input layouts of the reference (mnist_keras.py:50-57); channels_last is the default and the one
        # the fused sm_100a engine runs, channels_first trains on the torch-autograd backend
        shape = (1, 28, 28) if keras.backend.image_data_format() == "channels_first" else (28, 28, 1)
        return (x.reshape(len(x), *shape).astype("float32")) / 255

    return (prep(xtr), keras.utils.to_categorical(ytr, CLASSES)), (prep(xte), keras.utils.to_categorical(yte, CLASSES))


def build_model():
    return keras.Sequential([
        keras.layers.Conv2D(32, (3, 3), activation="relu",
                            input_shape=(1, 28, 28) if keras.backend.image_data_format() == "channels_first" else (28, 28, 1)),
        keras.layers.Conv2D(64, (3, 3), activation="relu"),
        keras.layers.MaxPooling2D((2, 2)),
        keras.layers.Dropout(0.25),
        keras.layers.Flatten(),
        keras.layers.Dense(128, activation="relu"),
        keras.layers.Dropout(0.5),
        keras.layers.Dense(CLASSES, activation="softmax"),
    ])


def main():
    metrics_sink.init(sync_tensorboard=True)          # local stand-in for the Gradient metrics sidecar
    model_dir, export_dir = hvd.model_paths()          # $PS_MODEL_PATH or ./models
    hvd.init()                                         # also pins this process to GPU <local_rank>
    world, me = hvd.size(), hvd.rank()

    epochs = int(os.environ.get("HVD_EXAMPLE_EPOCHS", math.ceil(BASE_EPOCHS / world)))   # fewer epochs with more GPUs
    (x_train, y_train), (x_test, y_test) = load_data()
    if me == 0:
        print(f"{len(x_train)} train / {len(x_test)} test samples, shape {x_train.shape[1:]}, {world} rank(s)")

    model = build_model()
    optimizer = hvd.DistributedOptimizer(keras.optimizers.Adadelta(1.0 * world))         # LR scales with ranks
    model.compile(loss=keras.losses.categorical_crossentropy, optimizer=optimizer, metrics=["accuracy"])

    callbacks = [hvd.callbacks.BroadcastGlobalVariablesCallback(0)]      # every rank starts from rank 0's state
    if me == 0:                                                          # only rank 0 writes files
        callbacks += [keras.callbacks.ModelCheckpoint(os.path.join(model_dir, "checkpoint-{epoch}.h5")),
                      keras.callbacks.TensorBoard(log_dir=os.path.join(model_dir, "eval"), update_freq="batch")]

    model.fit(x_train, y_train, batch_size=BATCH, epochs=epochs, callbacks=callbacks,
              verbose=int(me == 0), validation_data=(x_test, y_test))
    loss, accuracy = model.evaluate(x_test, y_test, verbose=0)

    if me == 0:
        final_path = os.path.join(model_dir, "keras-sample-model.h5")
        model.save(final_path)
        print(f"Model saved to {final_path}")
        keras.backend.set_learning_phase(0)
        served = keras.load_model(final_path)
        saved_model.export(served, os.path.join(export_dir, time.strftime("%Y%m%d-%H%M%S")),
                           inputs={"input": None}, outputs={"prob": None})
        print("acceptance:", metrics_sink.check("tensorflow:loss", "0.0..0.3", "mean"))

    print("Test loss:", loss)            # every rank reports its own evaluation, as the reference does
    print("Test accuracy:", accuracy)


if __name__ == "__main__":
    main()
