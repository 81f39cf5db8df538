"""LocalEstimator: single-process training on in-memory records (reference: zoo/.../examples/localEstimator/LenetLocalEstimator.scala
-- LeNet on MNIST with `LocalEstimator(model, criterion, optimMethod, validations, threadNum).fit(train, test, transformer, batch, epoch)`)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch.nn as nn

from zoo.pipeline.api.keras.optimizers import Adam
from zoo.pipeline.estimator import LocalEstimator

rng = np.random.RandomState(0)


def digits(n):
    """28x28 images whose class is the quadrant pattern of a blob (an MNIST-sized stand-in)."""
    y = rng.randint(0, 4, n)
    x = rng.rand(n, 1, 28, 28).astype(np.float32) * 0.3
    for i, c in enumerate(y):
        r, col = divmod(int(c), 2)
        x[i, 0, 4 + 12 * r: 12 + 12 * r, 4 + 12 * col: 12 + 12 * col] += 1.0
    return [(x[i], np.int64(y[i])) for i in range(n)]


lenet = nn.Sequential(nn.Conv2d(1, 6, 5), nn.Tanh(), nn.MaxPool2d(2), nn.Conv2d(6, 12, 5), nn.Tanh(), nn.MaxPool2d(2), nn.Flatten(),
                      nn.Linear(12 * 4 * 4, 100), nn.Tanh(), nn.Linear(100, 4))
est = LocalEstimator(lenet, nn.CrossEntropyLoss(), Adam(lr=2e-3), validations=["top1", "loss"], thread_num=4)
train, test = digits(2000), digits(400)
est.fit(train, test, batch_size=100, epoch=int(os.environ.get("EPOCHS", 3)))
print("validation:", {r.name: round(float(r.value), 4) for r in est.validate(test, batch_size=100)})
