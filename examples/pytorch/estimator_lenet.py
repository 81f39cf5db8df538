"""Estimator.from_torch with creator functions (LeNet on MNIST-shaped data).  `--workers 2` spawns two data-parallel
workers from this single process (one per GPU; gloo on CPU)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.orca.learn.pytorch import Estimator


class LeNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1, self.conv2 = nn.Conv2d(1, 20, 5), nn.Conv2d(20, 50, 5)
        self.fc1, self.fc2 = nn.Linear(800, 500), nn.Linear(500, 10)

    def forward(self, x):
        x = torch.max_pool2d(torch.relu(self.conv1(x)), 2)
        x = torch.max_pool2d(torch.relu(self.conv2(x)), 2)
        return torch.log_softmax(self.fc2(torch.relu(self.fc1(x.flatten(1)))), -1)


def data_creator(config, batch_size):
    rng = np.random.RandomState(0)
    y = rng.randint(0, 10, config["n"])
    x = rng.randn(config["n"], 1, 28, 28).astype(np.float32) * 0.1
    for k in range(10):                                               # class k lights up one image row
        x[y == k, 0, 2 * k + 4] += 1.0
    return x, y.astype(np.int64)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch-size", type=int, default=256)
    ap.add_argument("--workers", type=int, default=1)
    ap.add_argument("--n", type=int, default=4096)
    a = ap.parse_args()
    est = Estimator.from_torch(model_creator=lambda c: LeNet(), optimizer_creator=lambda m, c: torch.optim.Adam(m.parameters(), lr=c["lr"]),
                               loss_creator=lambda c: nn.NLLLoss(), metrics=["accuracy"], config={"lr": 1e-3, "n": a.n}, workers_per_node=a.workers)
    print(est.fit(data_creator, epochs=a.epochs, batch_size=a.batch_size))
    print(est.evaluate(data_creator, batch_size=a.batch_size))
