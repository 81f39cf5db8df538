"""Image classification with a TensorFlow model through TFNet (reference notebook: apps/tfnet/image_classification_inference.ipynb:
freeze a TF-Slim InceptionV1 with `export_tf`, `TFNet.from_export_folder`, read images with ImageSet, resize / normalise /
to-tensor, `predict`, map the arg-max to `imagenet_class_index.json`).

No TensorFlow and no network here: the frozen graph (frozen_inference_graph.pb + graph_meta.json, the reference's artefacts) is
written by this framework's `export_tf` from a small CNN with depthwise-separable blocks, then imported by its own GraphDef
reader and run on the GPU as a torch module."""
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "examples"))
import _common  # noqa: F401
import cv2
import numpy as np
import torch
import torch.nn as nn

from zoo.feature.common import ChainedPreprocessing
from zoo.feature.image import ImageCenterCrop, ImageChannelNormalize, ImageMatToTensor, ImageResize, ImageSet
from zoo.pipeline.api.net import TFNet
from zoo.util.tf import export_tf

# ## 1. A "pre-trained" model, frozen the way `export_tf` freezes a TF session
torch.manual_seed(0)
CLASSES = 20
cnn = nn.Sequential(nn.Conv2d(3, 16, 3, stride=2, padding=1), nn.BatchNorm2d(16), nn.ReLU6(),
                    nn.Conv2d(16, 16, 3, padding=1, groups=16), nn.BatchNorm2d(16), nn.ReLU6(), nn.Conv2d(16, 32, 1), nn.ReLU6(),
                    nn.AvgPool2d(8), nn.Flatten(), nn.Linear(32 * 14 * 14, CLASSES), nn.Softmax(-1)).eval()
folder = tempfile.mkdtemp()
export_tf(cnn, folder, torch.zeros(1, 3, 224, 224))
class_index = {str(i): ["n%08d" % i, "class_%d" % i] for i in range(CLASSES)}
with open(os.path.join(folder, "class_index.json"), "w") as fh:
    json.dump(class_index, fh)
print("export folder:", sorted(os.listdir(folder)))

# ## 2. Load it with TFNet
net = TFNet.from_export_folder(folder)

# ## 3. Images -> NHWC float tensors
rng = np.random.RandomState(0)
img_dir = tempfile.mkdtemp()
for i in range(4):
    cv2.imwrite(os.path.join(img_dir, "test%d.jpg" % i), rng.randint(0, 255, (300, 400, 3)).astype(np.uint8))
image_set = ImageSet.read(img_dir).transform(ChainedPreprocessing([ImageResize(256, 256), ImageCenterCrop(224, 224),
                                                                   ImageChannelNormalize(123.0, 117.0, 104.0, 58.0, 57.0, 57.0),
                                                                   ImageMatToTensor(format="NHWC")]))
batch = np.stack(image_set.get_image(to_chw=False))

# ## 4. Predict and decode
probs = np.asarray(net.predict(batch, batch_per_thread=4))
with open(os.path.join(folder, "class_index.json")) as fh:
    names = json.load(fh)
for feature, p in zip(image_set.get_image_features(), probs):
    top = np.argsort(-p)[:3]
    print(os.path.basename(feature["uri"]), "->", [(names[str(int(c))][1], round(float(p[c]), 3)) for c in top])
with torch.no_grad():
    want = cnn(torch.from_numpy(batch).permute(0, 3, 1, 2)).numpy()
print("classified %d images; max |TFNet - source model| = %.1e" % (len(probs), float(np.abs(probs - want).max())))
