"""Image similarity search (reference app: apps/image-similarity/image-similarity.ipynb -- real-estate photos: a scene classifier
fine-tuned with NNClassifier, then the pooled features of the network as embeddings, cosine similarity for "similar houses").
Images are generated (class = texture); embeddings come from the penultimate layer of the trained classifier."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.pipeline.api.keras.optimizers import Adam
from zoo.pipeline.estimator import Estimator
from zoo.feature.common import FeatureSet

rng = np.random.RandomState(0)
torch.manual_seed(0)


def texture(kind):
    yy, xx = np.mgrid[0:32, 0:32].astype(np.float32)
    base = [np.sin(xx / 2.0), np.sin(yy / 2.0), np.sin((xx + yy) / 3.0), ((xx // 4 + yy // 4) % 2) * 2 - 1][kind]
    return np.stack([base + 0.3 * rng.randn(32, 32).astype(np.float32)] * 3)


labels = rng.randint(0, 4, 800)
images = np.stack([texture(k) for k in labels]).astype(np.float32)
backbone = nn.Sequential(nn.Conv2d(3, 16, 5, stride=2), nn.ReLU(), nn.Conv2d(16, 32, 3, stride=2), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten())
net = nn.Sequential(backbone, nn.Linear(32, 4))
est = Estimator(net, optim_methods=Adam(lr=3e-3))
from zoo.common.triggers import MaxEpoch

est.train(FeatureSet.numpy(images[:640], labels[:640].astype(np.int64)), nn.CrossEntropyLoss(), end_trigger=MaxEpoch(int(os.environ.get("EPOCHS", 6))),
          batch_size=64)
net.eval()
dev = next(net.parameters()).device
with torch.no_grad():
    emb = nn.functional.normalize(backbone(torch.from_numpy(images).to(dev)), dim=1).cpu()
gallery, queries = emb[:640], emb[640:]
sim = queries @ gallery.t()
top = sim.topk(5, dim=1).indices.numpy()
hits = np.mean([np.mean(labels[:640][t] == labels[640 + i]) for i, t in enumerate(top)])
print("precision@5 of texture retrieval by cosine similarity of embeddings: %.3f" % hits)
