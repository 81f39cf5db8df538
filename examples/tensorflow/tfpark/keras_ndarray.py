"""tf.keras-style training on in-memory arrays through tfpark.KerasModel (reference example:
pyzoo/zoo/examples/tensorflow/tfpark/keras/keras_ndarray.py: a Sequential LeNet-ish model, `KerasModel(model)`,
`fit(x, y, batch_size, epochs, validation_data, distributed=True)`, `evaluate`, `predict`, `save_weights`)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import _common  # noqa: F401
import numpy as np
import torch

from zoo.pipeline.api.keras.layers import Convolution2D, Dense, Flatten, MaxPooling2D
from zoo.pipeline.api.keras.models import Sequential
from zoo.tfpark import KerasModel

torch.manual_seed(0)
rng = np.random.RandomState(0)
n, classes = int(os.environ.get("ROWS", 2048)), 10
protos = rng.rand(classes, 1, 28, 28).astype(np.float32)
y = rng.randint(0, classes, n).astype(np.int64)
x = protos[y] + 0.3 * rng.randn(n, 1, 28, 28).astype(np.float32)
split = n * 3 // 4

model = Sequential()
model.add(Convolution2D(6, 5, 5, activation="tanh", input_shape=(1, 28, 28)))
model.add(MaxPooling2D())
model.add(Convolution2D(12, 5, 5, activation="tanh"))
model.add(MaxPooling2D())
model.add(Flatten())
model.add(Dense(100, activation="tanh"))
model.add(Dense(classes, activation="softmax"))
model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])

keras_model = KerasModel(model)
keras_model.fit(x[:split], y[:split], batch_size=256, epochs=int(os.environ.get("EPOCHS", 4)), validation_data=(x[split:], y[split:]), distributed=True)
result = keras_model.evaluate(x[split:], y[split:], distributed=True, batch_per_thread=80)
print("evaluate:", result)
probs = np.asarray(keras_model.predict(x[split:split + 8], distributed=True, batch_per_thread=8))
print("predicted classes of 8 records:", probs.argmax(-1).tolist(), "labels:", y[split:split + 8].tolist())
path = os.path.join(tempfile.mkdtemp(), "model.h5")
keras_model.save_weights(path)
print("weights saved:", os.path.getsize(path) > 0)
