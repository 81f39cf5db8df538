"""ImageClassifier.predict_image_set on a folder of images (reference example: pyzoo/zoo/examples/imageclassification/predict.py:
`ImageClassifier.load_model(path)`, `ImageSet.read(folder)`, `predict_image_set`, `LabelOutput` -> top classes per image).
The images are generated here; the network is a small random-init CNN saved as a TorchScript archive."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import cv2
import numpy as np
import torch
import torch.nn as nn

from zoo.feature.image import ImageSet
from zoo.models.image.imageclassification import ImageClassifier

tmp = tempfile.mkdtemp()
rng = np.random.RandomState(0)
for i in range(6):
    cv2.imwrite(os.path.join(tmp, f"img{i}.jpg"), rng.randint(0, 255, (300, 400, 3)).astype(np.uint8))
torch.manual_seed(0)
net = nn.Sequential(nn.Conv2d(3, 8, 7, stride=4), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(8, 1000), nn.Softmax(-1)).eval()
path = os.path.join(tmp, "tiny-classifier.pt")
torch.jit.script(net).save(path)

model = ImageClassifier.load_model(path)
image_set = ImageSet.read(tmp)
output = model.predict_image_set(image_set)
for f in output.get_image_features():
    print(os.path.basename(f["uri"]), "->", [(c, round(float(p), 4)) for c, p in zip(f["clses"][:3], f["probs"][:3])])
print("predicted", len(output.get_image_features()), "images")
