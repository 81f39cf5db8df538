"""ImageNet-style training loop (reference examples: pyzoo/zoo/examples/inception/inception.py and the Scala TrainImageNet:
ImageSet -> random crop / flip / channel normalise -> SGD with warm-up + polynomial decay -> Top1/Top5 validation every epoch,
checkpoints).  A small Inception-like network on generated images so it runs anywhere; swap `build_net` and the folder for
the real thing."""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.common.triggers import EveryEpoch, MaxEpoch
from zoo.feature.common import FeatureSet
from zoo.pipeline.api.keras.optimizers import Poly, SequentialSchedule, Warmup
from zoo.pipeline.api.keras.optimizers import SGD
from zoo.pipeline.estimator import Estimator


class InceptionBlock(nn.Module):
    def __init__(self, cin, c1, c3, c5, cp):
        super().__init__()
        self.b1 = nn.Sequential(nn.Conv2d(cin, c1, 1), nn.ReLU())
        self.b3 = nn.Sequential(nn.Conv2d(cin, c3 // 2, 1), nn.ReLU(), nn.Conv2d(c3 // 2, c3, 3, padding=1), nn.ReLU())
        self.b5 = nn.Sequential(nn.Conv2d(cin, c5 // 2, 1), nn.ReLU(), nn.Conv2d(c5 // 2, c5, 5, padding=2), nn.ReLU())
        self.bp = nn.Sequential(nn.MaxPool2d(3, 1, 1), nn.Conv2d(cin, cp, 1), nn.ReLU())

    def forward(self, x):
        return torch.cat([self.b1(x), self.b3(x), self.b5(x), self.bp(x)], 1)


def build_net(classes):
    return nn.Sequential(nn.Conv2d(3, 16, 5, stride=2, padding=2), nn.ReLU(), nn.MaxPool2d(2), InceptionBlock(16, 8, 16, 8, 8),
                         InceptionBlock(40, 16, 24, 8, 8), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Dropout(0.2), nn.Linear(56, classes))


def make_data(n, classes, rng):
    y = rng.randint(0, classes, n)
    x = rng.randn(n, 3, 64, 64).astype(np.float32) * 0.5
    for i, c in enumerate(y):                       # class = position of a bright square
        r, col = divmod(int(c), 2)
        x[i, :, 8 + 24 * r: 28 + 24 * r, 8 + 24 * col: 28 + 24 * col] += 2.0
    return x, y.astype(np.int64)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=6)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--classes", type=int, default=4)
    a = ap.parse_args()
    rng = np.random.RandomState(0)
    xtr, ytr = make_data(1024, a.classes, rng)
    xva, yva = make_data(256, a.classes, rng)
    iters = a.epochs * (len(xtr) // a.batch_size)
    warm = max(1, iters // 10)
    sched = SequentialSchedule(len(xtr) // a.batch_size).add(Warmup(0.05 / warm), warm).add(Poly(0.5, iters - warm), iters - warm)
    est = Estimator(build_net(a.classes), optim_methods=SGD(learningrate=0.01, momentum=0.9, weightdecay=1e-4, leaningrate_schedule=sched),
                    model_dir=tempfile.mkdtemp())
    est.train(FeatureSet.numpy(xtr, ytr), nn.CrossEntropyLoss(), end_trigger=MaxEpoch(a.epochs), checkpoint_trigger=EveryEpoch(),
              validation_set=FeatureSet.numpy(xva, yva), validation_method=["top1", "loss"], batch_size=a.batch_size)
    res = est.evaluate(FeatureSet.numpy(xva, yva), ["top1", "top5" if a.classes > 5 else "loss"], batch_size=a.batch_size, criterion=nn.CrossEntropyLoss())
    print("validation:", {r.name: round(float(r.value), 4) for r in res})


if __name__ == "__main__":
    main()
