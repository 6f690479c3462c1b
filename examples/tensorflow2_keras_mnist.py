"""Equivalent of the reference's ``tensorflow2_keras_mnist.py`` on this framework: per-rank
dataset cache file, repeat/shuffle/batch pipeline, Adam with LR 0.001*size, the three hvd
callbacks in the reference's order, rank-0 checkpoint + TensorBoard, ``500 // size`` steps
x 24 epochs.  ``HVD_EXAMPLE_STEPS`` / ``HVD_EXAMPLE_EPOCHS`` shrink it for smoke tests.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras

model_dir, export_dir = hvd.model_paths()

hvd.init()   # pins the GPU for this local rank when there is one

(mnist_images, mnist_labels), _ = \
    keras.datasets.mnist.load_data(path='mnist-%d.npz' % hvd.rank())

dataset = keras.Dataset.from_tensor_slices(
    ((mnist_images[..., np.newaxis] / 255.0).astype(np.float32),
     mnist_labels.astype(np.int64))
)
dataset = dataset.repeat().shuffle(10000).batch(128)

mnist_model = keras.Sequential([
    keras.layers.Conv2D(32, [3, 3], activation='relu'),
    keras.layers.Conv2D(64, [3, 3], activation='relu'),
    keras.layers.MaxPooling2D(pool_size=(2, 2)),
    keras.layers.Dropout(0.25),
    keras.layers.Flatten(),
    keras.layers.Dense(128, activation='relu'),
    keras.layers.Dropout(0.5),
    keras.layers.Dense(10, activation='softmax')
])

opt = keras.optimizers.Adam(0.001 * hvd.size())
opt = hvd.DistributedOptimizer(opt)

mnist_model.compile(loss=keras.losses.SparseCategoricalCrossentropy(),
                    optimizer=opt,
                    metrics=['accuracy'],
                    experimental_run_tf_function=False)

callbacks = [
    hvd.callbacks.BroadcastGlobalVariablesCallback(0),
    # must come before any metrics-consuming callback
    hvd.callbacks.MetricAverageCallback(),
    hvd.callbacks.LearningRateWarmupCallback(warmup_epochs=3, verbose=1),
]

if hvd.rank() == 0:
    checkpoint_path = os.path.join(model_dir, 'checkpoint-{epoch}.h5')
    callbacks.append(keras.callbacks.ModelCheckpoint(checkpoint_path))
    callbacks.append(keras.callbacks.TensorBoard(log_dir=model_dir, update_freq='batch'))

verbose = 1 if hvd.rank() == 0 else 0

steps = int(os.environ.get("HVD_EXAMPLE_STEPS", 500 // hvd.size()))
epochs = int(os.environ.get("HVD_EXAMPLE_EPOCHS", 24))
mnist_model.fit(dataset, steps_per_epoch=steps, callbacks=callbacks, epochs=epochs, verbose=verbose)
