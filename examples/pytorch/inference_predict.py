"""Batch inference of a torchvision model through TorchNet (reference example: pyzoo/zoo/examples/pytorch/inference/predict.py:
an ImageFolder run through Resize / CenterCrop / Normalize, `TorchNet.from_pytorch(resnet18, [1, 3, 224, 224])`,
`net.predict(batch, distributed=True)`).  No pretrained weights or image folder in this image: the weights are fresh and the folder
is written first; `--image` points at a real folder."""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import cv2
import numpy as np
import torch
import torchvision

from zoo.common.nncontext import init_nncontext
from zoo.feature.common import ChainedPreprocessing
from zoo.feature.image import ImageCenterCrop, ImageChannelNormalize, ImageMatToTensor, ImageResize, ImageSet
from zoo.pipeline.api.net import TorchNet

ap = argparse.ArgumentParser()
ap.add_argument("--image", default=None, help="folder of images (or class sub-folders)")
ap.add_argument("--size", type=int, default=int(os.environ.get("IMAGE_SIZE", 128)))
a = ap.parse_args()
init_nncontext("Torch ResNet Prediction Example")
folder = a.image
if folder is None:
    folder = tempfile.mkdtemp(prefix="zoo_torch_pred_")
    rng = np.random.RandomState(0)
    for i in range(10):
        cv2.imwrite(os.path.join(folder, "img_%02d.jpg" % i), rng.randint(0, 255, (160 + 8 * i, 200, 3)).astype(np.uint8))
s = a.size
chain = ChainedPreprocessing([ImageResize(s + s // 8, s + s // 8), ImageCenterCrop(s, s),
                              ImageChannelNormalize(0.485 * 255, 0.456 * 255, 0.406 * 255, 0.229 * 255, 0.224 * 255, 0.225 * 255), ImageMatToTensor()])
images = ImageSet.read(folder).transform(chain)
batch = np.stack([f.get_tensor() for f in images.get_image_features()])
torch.manual_seed(0)
net = TorchNet.from_pytorch(torchvision.models.resnet18().eval(), [1, 3, s, s])
index = []
for k in range(0, len(batch), 8):                                   # batches of 8, as the reference's DataLoader
    out = net.predict(batch[k:k + 8], distributed=True)
    out = out.collect() if hasattr(out, "collect") else out
    index += [int(np.asarray(o).argmax()) for o in out]
print("predicted classes of %d images:" % len(index), index)
