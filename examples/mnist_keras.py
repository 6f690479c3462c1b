"""Equivalent of the reference's ``mnist_keras.py`` on this framework.

Same structure, same hyper-parameters (batch 128, ceil(12/size) epochs, Adadelta with LR
1.0*size, BroadcastGlobalVariables, rank-0 ModelCheckpoint + TensorBoard, fit with
validation, evaluate, rank-0 save / reload / export, every rank prints its score).
Run single-process (``python examples/mnist_keras.py``) or distributed
(``python -m horovod_distributed_example_b200.run -np 8 examples/mnist_keras.py``)
without code changes.  ``HVD_EXAMPLE_SAMPLES`` / ``HVD_EXAMPLE_EPOCHS`` shrink the run
for smoke tests.
"""
from __future__ import print_function

import math
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import horovod_distributed_example_b200 as hvd
from horovod_distributed_example_b200 import keras
from horovod_distributed_example_b200.keras import backend as K
from horovod_distributed_example_b200.keras.datasets import mnist
from horovod_distributed_example_b200.keras.layers import (Conv2D, Dense, Dropout, Flatten,
                                                           MaxPooling2D)
from horovod_distributed_example_b200.keras.models import Sequential, load_model
from horovod_distributed_example_b200.utils import metrics_sink as metrics
from horovod_distributed_example_b200.utils import saved_model

metrics.init(sync_tensorboard=True)

# Gradient path setup
model_dir, export_dir = hvd.model_paths()

# Initialize (also pins this process to GPU local_rank; no-op on a CPU host)
hvd.init()

batch_size = 128
num_classes = 10

# adjust number of epochs based on number of GPUs
epochs = int(os.environ.get("HVD_EXAMPLE_EPOCHS", int(math.ceil(12.0 / hvd.size()))))

img_rows, img_cols = 28, 28

(x_train, y_train), (x_test, y_test) = mnist.load_data()
n_small = int(os.environ.get("HVD_EXAMPLE_SAMPLES", 0))
if n_small:
    x_train, y_train = x_train[:n_small], y_train[:n_small]
    x_test, y_test = x_test[:max(n_small // 4, 256)], y_test[:max(n_small // 4, 256)]

if K.image_data_format() == 'channels_first':
    x_train = x_train.reshape(x_train.shape[0], 1, img_rows, img_cols)
    x_test = x_test.reshape(x_test.shape[0], 1, img_rows, img_cols)
    input_shape = (1, img_rows, img_cols)
else:
    x_train = x_train.reshape(x_train.shape[0], img_rows, img_cols, 1)
    x_test = x_test.reshape(x_test.shape[0], img_rows, img_cols, 1)
    input_shape = (img_rows, img_cols, 1)

x_train = x_train.astype('float32')
x_test = x_test.astype('float32')
x_train /= 255
x_test /= 255
print('x_train shape:', x_train.shape)
print(x_train.shape[0], 'train samples')
print(x_test.shape[0], 'test samples')

y_train = keras.utils.to_categorical(y_train, num_classes)
y_test = keras.utils.to_categorical(y_test, num_classes)

model = Sequential()
model.add(Conv2D(32, kernel_size=(3, 3), activation='relu', input_shape=input_shape))
model.add(Conv2D(64, (3, 3), activation='relu'))
model.add(MaxPooling2D(pool_size=(2, 2)))
model.add(Dropout(0.25))
model.add(Flatten())
model.add(Dense(128, activation='relu'))
model.add(Dropout(0.5))
model.add(Dense(num_classes, activation='softmax'))

# adjust learning rate based on number of GPUs, then wrap
opt = keras.optimizers.Adadelta(1.0 * hvd.size())
opt = hvd.DistributedOptimizer(opt)

model.compile(loss=keras.losses.categorical_crossentropy, optimizer=opt, metrics=['accuracy'])

callbacks = [
    # broadcast initial variable states from rank 0 to all other processes
    hvd.callbacks.BroadcastGlobalVariablesCallback(0),
]

# save checkpoints only on worker 0
if hvd.rank() == 0:
    checkpoint_path = os.path.join(model_dir, 'checkpoint-{epoch}.h5')
    logs_path = os.path.join(model_dir, 'eval')
    callbacks.append(keras.callbacks.ModelCheckpoint(checkpoint_path))
    callbacks.append(keras.callbacks.TensorBoard(log_dir=logs_path, update_freq='batch'))

model.fit(x_train, y_train,
          batch_size=batch_size,
          callbacks=callbacks,
          epochs=epochs,
          verbose=1 if hvd.rank() == 0 else 0,
          validation_data=(x_test, y_test))
score = model.evaluate(x_test, y_test, verbose=0)

if hvd.rank() == 0:
    exported_model_path = os.path.join(model_dir, 'keras-sample-model.h5')
    model.save(exported_model_path)
    print("Model saved to {}".format(exported_model_path))

    K.set_learning_phase(0)
    new_model = load_model(exported_model_path)

    # serving export: input -> prob signature under <export_dir>/<timestamp>
    saved_model.export(new_model, os.path.join(export_dir, time.strftime("%Y%m%d-%H%M%S")),
                       inputs={'input': None}, outputs={'prob': None})
    print("CI check:", metrics.check("tensorflow:loss", "0.0..0.3", "mean"))

print('Test loss:', score[0])
print('Test accuracy:', score[1])
