"""3-D image augmentation (reference app: apps/image-augmentation-3d/image-augmentation-3d.ipynb -- a CT-like volume through Crop3D,
RandomCrop3D, CenterCrop3D, Rotate3D and AffineTransform3D)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.feature.image3d.transformation import AffineTransform3D, CenterCrop3D, Crop3D, RandomCrop3D, Rotate3D
from zoo.feature.image.imageset import ImageFeature

np.random.seed(0)
zz, yy, xx = np.mgrid[0:32, 0:48, 0:48].astype(np.float32)
volume = np.exp(-(((zz - 16) / 6) ** 2 + ((yy - 20) / 10) ** 2 + ((xx - 30) / 4) ** 2))         # an off-centre ellipsoid


def run(op):
    f = op.apply(ImageFeature(imageTensor=volume.copy()))
    return np.asarray(f["imageTensor"])


def centroid(v):
    idx = np.indices(v.shape).reshape(3, -1)
    w = v.reshape(-1)
    return np.round((idx * w).sum(1) / w.sum(), 1).tolist()


print("source          ", volume.shape, "centroid", centroid(volume))
for name, op in [("crop", Crop3D([4, 8, 8], [20, 32, 32])), ("random crop", RandomCrop3D(20, 32, 32)), ("center crop", CenterCrop3D(20, 32, 32)),
                 ("rotate 30 deg", Rotate3D([0.0, 0.0, np.pi / 6])),
                 ("affine (scale)", AffineTransform3D(np.diag([1.0, 0.8, 1.25]), np.zeros(3), "clamp"))]:
    out = run(op)
    print(f"{name:16s}", out.shape, "centroid", centroid(out))
