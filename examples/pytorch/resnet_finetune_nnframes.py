"""Fine-tune a torchvision ResNet through NNFrames (the reference's dogs-vs-cats path: TorchNet + TorchCriterion +
NNClassifier on a DataFrame of images), on synthetic two-class images."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import pandas as pd
import torch.nn as nn
import torchvision

from zoo.common.triggers import EveryEpoch
from zoo.feature.common import ChainedPreprocessing
from zoo.feature.image import ImageCenterCrop, ImageChannelNormalize, ImageFeatureToTensor, ImageMatToTensor, ImageResize, RowToImageFeature
from zoo.pipeline.api.keras.optimizers import Adam
from zoo.pipeline.api.net import TorchCriterion, TorchNet
from zoo.pipeline.nnframes import NNClassifier

ap = argparse.ArgumentParser()
ap.add_argument("--epochs", type=int, default=1)
ap.add_argument("--images", type=int, default=64)
ap.add_argument("--size", type=int, default=64)
a = ap.parse_args()
rng = np.random.RandomState(0)
rows = []
for i in range(a.images):
    label = i % 2
    img = rng.randint(0, 80, (a.size + 16, a.size + 16, 3), dtype=np.uint8)
    img[:, :, label * 2] += 120                                        # class 0 is blue-ish, class 1 red-ish (BGR)
    rows.append({"image": {"origin": f"mem:{i}", "height": img.shape[0], "width": img.shape[1], "nChannels": 3, "mode": 16, "data": img},
                 "label": float(label + 1)})
df = pd.DataFrame(rows)
net = torchvision.models.resnet18(num_classes=2)
model = TorchNet.from_pytorch(nn.Sequential(net, nn.LogSoftmax(dim=-1)), [2, 3, a.size, a.size])
crit = TorchCriterion.from_pytorch(lambda out, y: nn.functional.nll_loss(out, (y.reshape(-1) - 1).long()))
chain = ChainedPreprocessing([RowToImageFeature(), ImageResize(a.size + 8, a.size + 8), ImageCenterCrop(a.size, a.size),
                              ImageChannelNormalize(123.0, 117.0, 104.0, 58.0, 57.0, 57.0), ImageMatToTensor(), ImageFeatureToTensor()])
clf = NNClassifier(model, crit, chain).setBatchSize(16).setMaxEpoch(a.epochs).setFeaturesCol("image").setOptimMethod(Adam(lr=1e-3)) \
    .setValidation(EveryEpoch(), df, ["accuracy"], 16)
out = clf.fit(df).transform(df)
print("training-set accuracy:", float((out["prediction"] == df["label"]).mean()))
