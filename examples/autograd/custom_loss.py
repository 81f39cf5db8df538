"""autograd: a custom loss and a Lambda layer written with Variable algebra (reference examples:
pyzoo/zoo/examples/autograd/custom.py and customloss.py -- mean absolute error as CustomLoss, `add_one` as a Lambda)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

import zoo.pipeline.api.autograd as A
from zoo.pipeline.api.keras.layers import Dense
from zoo.pipeline.api.keras.models import Sequential


def mean_absolute_error(y_true, y_pred):
    return A.mean(A.abs(y_true - y_pred), axis=1)


def add_one(x):
    return x + 1.0


rng = np.random.RandomState(0)
x = rng.uniform(0, 1, (1000, 2)).astype(np.float32)
y = ((2 * x).sum(1) + 0.4).reshape(1000, 1).astype(np.float32)
model = Sequential()
model.add(Dense(1, input_shape=(2,)))
model.add(A.Lambda(function=add_one))
model.compile(optimizer="adam", loss=A.CustomLoss(mean_absolute_error, y_pred_shape=(1,)))
before = model.evaluate(x, y, batch_size=500)[0]
model.fit(x, y, batch_size=32, nb_epoch=int(os.environ.get("EPOCHS", 20)))
after = model.evaluate(x, y, batch_size=500)[0]
print("mae before %.4f after %.4f" % (before, after))
w = model.get_weights()
print("weights", np.round(w[0], 2).tolist(), "bias", np.round(w[1], 2).tolist(), "(target: 2, 2 and 0.4 - 1)")
