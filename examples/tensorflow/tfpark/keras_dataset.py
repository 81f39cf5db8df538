"""tfpark.KerasModel over a TFDataset and over ndarrays (reference examples: pyzoo/zoo/examples/tensorflow/tfpark/keras/
keras_dataset.py and keras_ndarray.py: compile a Keras model, `KerasModel(model)`, `fit(dataset, epochs)` /
`fit(x, y, batch_size, distributed=True)`, `evaluate`, `predict`, `save_weights`)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import _common  # noqa: F401
import numpy as np
import torch

from zoo.pipeline.api.keras.layers import Dense, Dropout
from zoo.pipeline.api.keras.models import Sequential
from zoo.tfpark import KerasModel, TFDataset

torch.manual_seed(0)
rng = np.random.RandomState(2)
x = rng.randn(2000, 16).astype(np.float32)
y = (np.abs(x[:, :4]).sum(1) > 3.2).astype(np.int64)

net = Sequential()
net.add(Dense(64, activation="relu", input_shape=(16,)))
net.add(Dropout(0.1))
net.add(Dense(2, activation="softmax"))
net.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
model = KerasModel(net)

epochs = int(os.environ.get("EPOCHS", 20))
model.fit(TFDataset.from_ndarrays((x[:1600], y[:1600]), batch_size=100, val_tensors=(x[1600:], y[1600:])), epochs=epochs)   # dataset route
model.fit(x[:1600], y[:1600], batch_size=100, epochs=max(epochs // 4, 1), distributed=True)                                 # ndarray route
loss, acc = model.evaluate(x[1600:], y[1600:], batch_size=200)
path = os.path.join(tempfile.mkdtemp(), "weights.pt")
model.save_weights(path)
print("held-out loss %.3f accuracy %.3f; predictions %s; weights in %s" % (loss, acc, model.predict(x[1600:1604]).argmax(1).tolist(), path))
