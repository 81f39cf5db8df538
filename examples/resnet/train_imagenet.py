"""ResNet training recipe (reference example: zoo/.../examples/resnet/TrainImageNet.scala + README: ResNet-50 shortcut type B,
SGD momentum 0.9, weight decay 1e-4 (not on BN / bias), label smoothing-free cross entropy, learning-rate warm-up for
`--warmupEpoch` epochs from `--learningRate` up to `--maxLr`, then polynomial decay (power 2) over the remaining epochs, global
batch 8192, Top1 / Top5 / loss validation every epoch, checkpoints under `--cache`).

The same recipe on this framework's engine: bf16 autocast + channels_last through cuDNN, data-parallel over the visible GPUs
when launched with `python -m torch.distributed.run`.  With `-f / --folder` the data are the reference's ImageNet SequenceFiles
(`<folder>/train`, `<folder>/val`, read with `ImageSet.read_sequence_files`, random crop + flip for training, centre crop for
validation, channel normalisation: examples/resnet/Utils.scala:118-140); without it images are generated (no ImageNet in the
build environment).  `--depth` chooses the torchvision ResNet."""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn
import torchvision

from zoo.common.triggers import EveryEpoch, MaxEpoch
from zoo.feature.common import FeatureSet
from zoo.pipeline.api.keras.optimizers import SGD, Poly, SequentialSchedule, Warmup
from zoo.pipeline.estimator import Estimator

ap = argparse.ArgumentParser()
ap.add_argument("--batchSize", type=int, default=64)
ap.add_argument("--nEpochs", type=int, default=int(os.environ.get("EPOCHS", 4)))
ap.add_argument("--learningRate", type=float, default=0.02)
ap.add_argument("--warmupEpoch", type=int, default=1)
ap.add_argument("--maxLr", type=float, default=0.08)
ap.add_argument("--cache", default=None)
ap.add_argument("--depth", type=int, default=18, choices=[18, 34, 50, 101, 152])
ap.add_argument("--classes", type=int, default=4)
ap.add_argument("--images", type=int, default=int(os.environ.get("IMAGES", 512)))
ap.add_argument("--size", type=int, default=64)
ap.add_argument("-f", "--folder", default=None, help="folder with train/ and val/ SequenceFiles (labels 1..classes)")
a = ap.parse_args()

rng = np.random.RandomState(0)


def make(n):
    y = rng.randint(0, a.classes, n)
    x = (rng.randn(n, 3, a.size, a.size) * 0.4).astype(np.float32)
    h = a.size // 2
    for i, c in enumerate(y):                               # class = which quadrant is bright
        r, col = divmod(int(c) % 4, 2)
        x[i, :, r * h:(r + 1) * h, col * h:(col + 1) * h] += 1.5
    return x, y.astype(np.int64)


def from_sequence_files(folder, train):
    from zoo.feature.common import ChainedPreprocessing
    from zoo.feature.image import ImageCenterCrop, ImageChannelNormalize, ImageHFlip, ImageMatToTensor, ImageRandomCrop, ImageRandomPreprocessing, ImageSet

    crop = [ImageRandomCrop(a.size, a.size), ImageRandomPreprocessing(ImageHFlip(), 0.5)] if train else [ImageCenterCrop(a.size, a.size)]
    chain = ChainedPreprocessing(crop + [ImageChannelNormalize(104.0, 117.0, 123.0, 57.375, 57.12, 58.395), ImageMatToTensor()])
    feats = ImageSet.read_sequence_files(folder, class_num=a.classes).transform(chain).get_image_features()
    return np.stack([f.get_tensor() for f in feats]).astype(np.float32), np.asarray([int(f["label"][0]) - 1 for f in feats], np.int64)


if a.folder:
    xtr, ytr = from_sequence_files(os.path.join(a.folder, "train"), True)
    xva, yva = from_sequence_files(os.path.join(a.folder, "val"), False)
else:
    xtr, ytr = make(a.images)
    xva, yva = make(a.images // 4)
torch.manual_seed(0)
net = getattr(torchvision.models, f"resnet{a.depth}")(num_classes=a.classes)
for m in net.modules():                                      # the reference's init: last BN gamma of every residual branch = 0
    if isinstance(m, (torchvision.models.resnet.BasicBlock,)):
        nn.init.zeros_(m.bn2.weight)
    elif isinstance(m, (torchvision.models.resnet.Bottleneck,)):
        nn.init.zeros_(m.bn3.weight)

per_epoch = max(len(xtr) // a.batchSize, 1)
warm_iters = a.warmupEpoch * per_epoch
total = a.nEpochs * per_epoch
schedule = SequentialSchedule(per_epoch).add(Warmup((a.maxLr - a.learningRate) / max(warm_iters, 1)), warm_iters) \
    .add(Poly(2.0, max(total - warm_iters, 1)), max(total - warm_iters, 1))
optim = SGD(learningrate=a.learningRate, momentum=0.9, dampening=0.0, weightdecay=1e-4, leaningrate_schedule=schedule)
est = Estimator(net, optim_methods=optim, model_dir=a.cache or tempfile.mkdtemp())
est.train(FeatureSet.numpy(xtr, ytr), nn.CrossEntropyLoss(), end_trigger=MaxEpoch(a.nEpochs), checkpoint_trigger=EveryEpoch(),
          validation_set=FeatureSet.numpy(xva, yva), validation_method=["top1", "loss"], batch_size=a.batchSize)
res = est.evaluate(FeatureSet.numpy(xva, yva), ["top1", "loss"], batch_size=a.batchSize, criterion=nn.CrossEntropyLoss())
print("validation:", {r.name: round(float(r.value), 4) for r in res})
