"""Image generation with a PyTorch generator served through InferenceModel (reference app: apps/pytorch/face_generation.ipynb -- a
pretrained PGAN from torch hub wrapped in TorchNet and run on noise vectors; no network here, so the generator is a random-init
DCGAN of the same interface).  Shows the path: torch module -> TorchScript -> InferenceModel pool -> batched predict -> image grid."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.pipeline.inference import InferenceModel

torch.manual_seed(0)
generator = nn.Sequential(nn.Unflatten(1, (64, 1, 1)), nn.ConvTranspose2d(64, 64, 4), nn.BatchNorm2d(64), nn.ReLU(),
                          nn.ConvTranspose2d(64, 32, 4, 2, 1), nn.BatchNorm2d(32), nn.ReLU(), nn.ConvTranspose2d(32, 16, 4, 2, 1), nn.BatchNorm2d(16),
                          nn.ReLU(), nn.ConvTranspose2d(16, 3, 4, 2, 1), nn.Tanh()).eval()
path = os.path.join(tempfile.mkdtemp(), "generator.pt")
torch.jit.trace(generator, torch.randn(2, 64)).save(path)
model = InferenceModel(supported_concurrent_num=2)
model.load_torch(path) if hasattr(model, "load_torch") else model.load(path)
noise = np.random.RandomState(0).randn(16, 64).astype(np.float32)
imgs = np.asarray(model.predict(noise))
grid = ((imgs.reshape(4, 4, 3, 32, 32).transpose(0, 3, 1, 4, 2).reshape(128, 128, 3) + 1) * 127.5).astype(np.uint8)
out = os.path.join(os.path.dirname(path), "faces.png")
try:
    import cv2

    cv2.imwrite(out, grid[:, :, ::-1])
except ImportError:
    np.save(out + ".npy", grid)
print("generated", imgs.shape, "-> grid", grid.shape, "written to", out)
