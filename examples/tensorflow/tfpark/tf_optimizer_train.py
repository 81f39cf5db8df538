"""TFOptimizer / TFPredictor (reference examples: pyzoo/zoo/examples/tensorflow/tfpark/tf_optimizer/train.py and evaluate.py:
LeNet on MNIST, `TFOptimizer.from_loss(loss, Adam(1e-3))`, `optimizer.optimize(end_trigger=MaxEpoch(5))`, then a
`TFPredictor` pass to compute accuracy).  The graph is a torch module here; data is MNIST-shaped and synthetic (no network)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.common.triggers import MaxEpoch
from zoo.pipeline.api.keras.optimizers import Adam
from zoo.tfpark import TFDataset, TFOptimizer, TFPredictor

torch.manual_seed(0)
rng = np.random.RandomState(0)
N, CLASSES = int(os.environ.get("ROWS", 4096)), 10
protos = rng.rand(CLASSES, 1, 28, 28).astype(np.float32)                  # one blurry prototype per digit class
labels = rng.randint(0, CLASSES, N).astype(np.int64)
images = protos[labels] + 0.3 * rng.randn(N, 1, 28, 28).astype(np.float32)
split = N * 3 // 4

lenet = nn.Sequential(nn.Conv2d(1, 6, 5), nn.Tanh(), nn.MaxPool2d(2), nn.Conv2d(6, 12, 5), nn.Tanh(), nn.MaxPool2d(2), nn.Flatten(),
                      nn.Linear(12 * 4 * 4, 100), nn.Tanh(), nn.Linear(100, CLASSES))
dataset = TFDataset.from_ndarrays((images[:split], labels[:split]), batch_size=256, val_tensors=(images[split:], labels[split:]))
optimizer = TFOptimizer.from_loss(nn.CrossEntropyLoss(), Adam(lr=1e-3), model=lenet, dataset=dataset, val_method=["accuracy"])
optimizer.set_gradient_clipping_by_l2_norm(10.0)
optimizer.optimize(end_trigger=MaxEpoch(int(os.environ.get("EPOCHS", 5))))

predictions = TFPredictor.from_outputs(lenet, dataset=TFDataset.from_ndarrays(images[split:], batch_per_thread=256)).predict()
acc = float((np.stack(predictions).argmax(1) == labels[split:]).mean())
print("held-out accuracy %.3f over %d records" % (acc, len(predictions)))
