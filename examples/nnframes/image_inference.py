"""NNFrames image inference (reference example: pyzoo/zoo/examples/nnframes/imageInference/ImageInferenceExample.py:
`NNImageReader.readImages(folder)`, a preprocessing chain ending in `ImageFeatureToTensor`, `NNClassifierModel(model, chain)
.setFeaturesCol("image").setBatchSize(n)`, `transform(df)` adds a `prediction` column)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import cv2
import numpy as np
import torch
import torch.nn as nn

from zoo.feature.common import ChainedPreprocessing
from zoo.feature.image import ImageCenterCrop, ImageChannelNormalize, ImageFeatureToTensor, ImageMatToTensor, ImageResize, RowToImageFeature
from zoo.pipeline.nnframes import NNClassifierModel, NNImageReader

folder = tempfile.mkdtemp()
rng = np.random.RandomState(0)
for i in range(8):
    img = rng.randint(0, 60, (120, 160, 3)).astype(np.uint8)
    img[:, :, (i % 2) * 2] += 150                    # even files are blue-ish, odd files red-ish
    cv2.imwrite(os.path.join(folder, "img%d.png" % i), img)

torch.manual_seed(0)
net = nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(3, 2), nn.LogSoftmax(-1)).eval()
with torch.no_grad():                                   # a colour detector: class 1 = more blue, class 2 = more red
    net[2].weight.copy_(torch.tensor([[1.0, 0.0, -1.0], [-1.0, 0.0, 1.0]]))
    net[2].bias.zero_()
chain = ChainedPreprocessing([RowToImageFeature(), ImageResize(72, 72), ImageCenterCrop(64, 64), ImageChannelNormalize(100.0, 100.0, 100.0),
                              ImageMatToTensor(), ImageFeatureToTensor()])
df = NNImageReader.readImages(folder, resizeH=96, resizeW=96)
model = NNClassifierModel(net, chain).setFeaturesCol("image").setBatchSize(4)
out = model.transform(df).sort_values("image", key=lambda c: c.map(lambda r: r["origin"]))
for _, row in out.iterrows():
    print(os.path.basename(row["image"]["origin"]), "-> class", int(row["prediction"]))
print("images classified:", len(out), "distinct classes:", sorted(set(int(p) for p in out["prediction"])))
