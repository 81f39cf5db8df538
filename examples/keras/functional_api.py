"""Keras-style functional API: two inputs, a shared layer, autograd ops, a custom loss."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

import zoo.pipeline.api.autograd as A
from zoo.pipeline.api.keras.engine.topology import Model
from zoo.pipeline.api.keras.layers import Dense, Input, merge

a, b = Input(shape=(8,)), Input(shape=(8,))
shared = Dense(16, activation="relu")
h = merge([shared(a), shared(b)], mode="concat")
h = h * 0.5 + A.abs(h) * 0.5                                             # == relu(h), written with Variable algebra
out = Dense(1)(h)
model = Model([a, b], out)
model.summary()
loss = A.CustomLoss(lambda y_true, y_pred: A.mean(A.square(y_true - y_pred), axis=1), y_pred_shape=(1,))
model.compile(optimizer="adam", loss=loss)
rng = np.random.RandomState(0)
xa, xb = rng.randn(2000, 8).astype(np.float32), rng.randn(2000, 8).astype(np.float32)
y = (xa.sum(1, keepdims=True) - xb.sum(1, keepdims=True)).astype(np.float32)
print("mse before:", model.evaluate([xa, xb], y, batch_size=500)[0])
model.fit([xa, xb], y, batch_size=100, nb_epoch=10)
print("mse after :", model.evaluate([xa, xb], y, batch_size=500)[0])
