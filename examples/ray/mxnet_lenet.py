"""LeNet on MNIST-shaped data with `MXNetTrainer` (reference example: pyzoo/zoo/examples/ray/mxnet/lenet_mnist.py -- creator
functions for data / model / loss / metrics shipped to Ray actors, `kvstore=dist_sync`).  The trainer keeps that shape; the workers
are torch processes of the local pool and the synchronous parameter server is the gradient all-reduce.  Synthetic digits (no
network): class k is a bright k-th stripe on noise."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.ray.mxnet import MXNetTrainer
from zoo.ray.mxnet.utils import create_trainer_config


def get_data(config, batch_size):
    rng = np.random.RandomState(config.get("seed", 42))
    n = int(config.get("samples", 2048))
    y = rng.randint(0, 10, n)
    x = rng.rand(n, 1, 28, 28).astype(np.float32) * 0.3
    for i, k in enumerate(y):
        x[i, 0, 2 * k + 3:2 * k + 6] += 0.7
    return x, y.astype(np.int64)


def get_model(config):
    return nn.Sequential(nn.Conv2d(1, 20, 5), nn.Tanh(), nn.MaxPool2d(2), nn.Conv2d(20, 50, 5), nn.Tanh(), nn.MaxPool2d(2), nn.Flatten(),
                         nn.Linear(800, 500), nn.Tanh(), nn.Linear(500, 10))


def get_loss(config):
    return nn.CrossEntropyLoss()


def get_metrics(config):
    return ["accuracy"]


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("-n", "--num_workers", type=int, default=2)
    ap.add_argument("-b", "--batch_size", type=int, default=64, help="per worker, as in the reference")
    ap.add_argument("-e", "--epochs", type=int, default=int(os.environ.get("EPOCHS", 3)))
    ap.add_argument("-l", "--learning_rate", type=float, default=0.05)
    args = ap.parse_args()
    config = create_trainer_config(batch_size=args.batch_size, optimizer="sgd", optimizer_params={"learning_rate": args.learning_rate},
                                   log_interval=10, seed=42, extra_config={"lr": args.learning_rate})
    trainer = MXNetTrainer(config, get_data, get_model, get_loss, get_metrics, num_workers=args.num_workers)
    stats = trainer.train(nb_epoch=args.epochs)
    print("per-worker stats:", [{k: (round(v, 4) if isinstance(v, float) else v) for k, v in s.items() if k in ("iterations", "Loss", "loss")} for s in stats])
    x, y = get_data(dict(config, seed=7, samples=512), 0)
    with torch.no_grad():
        acc = float((trainer.est.get_model()(torch.from_numpy(x)).argmax(-1).numpy() == y).mean())
    print("held-out accuracy after %d epochs on %d workers: %.3f" % (args.epochs, args.num_workers, acc))
    trainer.shutdown()
