"""3-D image transforms (reference: pyzoo/zoo/feature/image3d/transformation.py:29-120, zoo/.../feature/image3d/*.scala):
Crop3D / RandomCrop3D / CenterCrop3D, Rotate3D (yaw, pitch, roll), AffineTransform3D (trilinear warp with clamp or
zero padding).  Volumes are (depth, height, width[, channels]) ndarrays under the feature's `imageTensor` key, or
plain ndarrays."""
from __future__ import annotations

import numpy as np

from ..common import Preprocessing


class ImagePreprocessing3D(Preprocessing):
    def __init__(self, bigdl_type="float", *args):
        pass

    def apply(self, f):
        if isinstance(f, dict):
            f["imageTensor"] = self.transform_volume(np.asarray(f["imageTensor"], dtype=np.float32))
            return f
        return self.transform_volume(np.asarray(f, dtype=np.float32))

    def transform_volume(self, v):  # pragma: no cover - interface
        return v


class Crop3D(ImagePreprocessing3D):
    def __init__(self, start, patch_size, bigdl_type="float"):
        self.start, self.patch = [int(s) for s in start], [int(s) for s in patch_size]

    def transform_volume(self, v):
        (z, y, x), (d, h, w) = self.start, self.patch
        if z + d > v.shape[0] or y + h > v.shape[1] or x + w > v.shape[2]:
            raise ValueError("crop exceeds the volume")
        return v[z:z + d, y:y + h, x:x + w]


class RandomCrop3D(ImagePreprocessing3D):
    def __init__(self, crop_depth, crop_height, crop_width, bigdl_type="float"):
        self.size = (int(crop_depth), int(crop_height), int(crop_width))

    def transform_volume(self, v):
        st = [np.random.randint(0, v.shape[i] - self.size[i] + 1) for i in range(3)]
        return Crop3D(st, self.size).transform_volume(v)


class CenterCrop3D(ImagePreprocessing3D):
    def __init__(self, crop_depth, crop_height, crop_width, bigdl_type="float"):
        self.size = (int(crop_depth), int(crop_height), int(crop_width))

    def transform_volume(self, v):
        st = [(v.shape[i] - self.size[i]) // 2 for i in range(3)]
        return Crop3D(st, self.size).transform_volume(v)


class AffineTransform3D(ImagePreprocessing3D):
    """dst(p) = src(A (p - c) + c + t) with trilinear interpolation; `clamp_mode` "clamp" replicates the border,
    "padding" uses `pad_val` outside."""

    def __init__(self, affine_mat, translation=np.zeros(3), clamp_mode="clamp", pad_val=0.0, bigdl_type="float"):
        self.A, self.t = np.asarray(affine_mat, dtype=np.float64).reshape(3, 3), np.asarray(translation, dtype=np.float64).reshape(3)
        if clamp_mode not in ("clamp", "padding"):
            raise ValueError("clamp_mode must be 'clamp' or 'padding'")
        self.clamp_mode, self.pad_val = clamp_mode, float(pad_val)

    def transform_volume(self, v):
        from scipy.ndimage import map_coordinates

        shape = v.shape[:3]
        c = (np.asarray(shape, dtype=np.float64) - 1) / 2.0
        grid = np.stack(np.meshgrid(*[np.arange(s, dtype=np.float64) for s in shape], indexing="ij"), 0).reshape(3, -1)
        src = self.A @ (grid - c[:, None]) + c[:, None] + self.t[:, None]
        mode, cval = ("nearest", 0.0) if self.clamp_mode == "clamp" else ("constant", self.pad_val)

        def warp(vol):
            return map_coordinates(vol, src, order=1, mode=mode, cval=cval).reshape(shape)

        if v.ndim == 4:
            return np.stack([warp(v[..., ch]) for ch in range(v.shape[3])], -1).astype(np.float32)
        return warp(v).astype(np.float32)


class Rotate3D(AffineTransform3D):
    """Rotation by (yaw, pitch, roll) radians about the volume centre."""

    def __init__(self, rotation_angles, bigdl_type="float"):
        yaw, pitch, roll = [float(a) for a in rotation_angles]
        cy, sy, cp, sp, cr, sr = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch), np.cos(roll), np.sin(roll)
        rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
        ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
        rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
        super().__init__(rz @ ry @ rx)
