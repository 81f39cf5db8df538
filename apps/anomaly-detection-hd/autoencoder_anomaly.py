"""Anomaly detection on high-dimensional data with an autoencoder (reference app: apps/anomaly-detection-hd/autoencoder-zoo.ipynb --
heartbeat / sensor vectors, Dense autoencoder trained on normal data with the Keras-style API, reconstruction error as the score)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.pipeline.api.keras.layers import Dense
from zoo.pipeline.api.keras.models import Sequential

rng = np.random.RandomState(0)
d, k = 48, 4
basis = rng.randn(k, d).astype(np.float32)
normal = (rng.randn(6000, k).astype(np.float32) @ basis) + 0.05 * rng.randn(6000, d).astype(np.float32)     # lives near a 4-D subspace
outliers = rng.randn(60, d).astype(np.float32) * 1.5                                                          # does not
mu, sd = normal.mean(0), normal.std(0)
norm = lambda a: (a - mu) / sd

model = Sequential()
model.add(Dense(24, activation="relu", input_shape=(d,)))
model.add(Dense(k))
model.add(Dense(24, activation="relu"))
model.add(Dense(d))
model.compile(optimizer="adam", loss="mse")
model.fit(norm(normal[:5000]), norm(normal[:5000]), batch_size=128, nb_epoch=int(os.environ.get("EPOCHS", 10)))
test = np.concatenate([norm(normal[5000:]), norm(outliers)])
truth = np.concatenate([np.zeros(1000, bool), np.ones(60, bool)])
err = ((np.asarray(model.predict(test, batch_size=256)) - test) ** 2).mean(1)
thr = np.sort(err)[-60]
flagged = err >= thr
print("top-60 reconstruction errors: %d of 60 injected outliers found (precision %.2f)" % ((flagged & truth).sum(), (flagged & truth).sum() / flagged.sum()))
