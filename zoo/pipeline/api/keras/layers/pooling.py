"""Pooling layers (reference: layers/pooling.py).  1-D layers take (batch, steps, dim)."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .base import _FnLayer


def _same_pad(x, k, s, nd, value):
    """Keras 'same' padding for pooling (pad so that out = ceil(in / stride))."""
    pads = []
    for d in range(nd):
        n = x.shape[-1 - d]
        kk, ss = k[-1 - d], s[-1 - d]
        total = max((((n + ss - 1) // ss) - 1) * ss + kk - n, 0)
        pads += [total // 2, total - total // 2]
    return F.pad(x, pads, value=value)


class _Pool(_FnLayer):
    nd, kind = 2, "max"

    def __init__(self, pool_size, strides, border_mode, dim_ordering, input_shape, **kw):
        super().__init__(input_shape, **kw)
        if border_mode not in ("valid", "same"):
            raise ValueError(f"border_mode must be 'valid' or 'same', got {border_mode}")
        self.pool_size = tuple(pool_size)
        self.strides = tuple(strides) if strides is not None else self.pool_size
        self.border_mode, self.dim_ordering = border_mode, dim_ordering

    def fn(self, x):
        x = x.float()
        if self.dim_ordering == "tf":
            x = x.movedim(-1, 1)
        if self.border_mode == "same":
            x = _same_pad(x, self.pool_size, self.strides, self.nd, float("-inf") if self.kind == "max" else 0.0)
        f = {("max", 1): F.max_pool1d, ("max", 2): F.max_pool2d, ("max", 3): F.max_pool3d,
             ("avg", 1): F.avg_pool1d, ("avg", 2): F.avg_pool2d, ("avg", 3): F.avg_pool3d}[(self.kind, self.nd)]
        y = f(x, self.pool_size, self.strides)
        return y.movedim(1, -1) if self.dim_ordering == "tf" else y


class MaxPooling1D(_Pool):
    nd, kind = 1, "max"

    def __init__(self, pool_length=2, stride=None, border_mode="valid", input_shape=None, pad=0, **kw):
        super().__init__((int(pool_length),), None if stride in (None, -1) else (int(stride),), border_mode, "tf", input_shape, **kw)


class AveragePooling1D(_Pool):
    nd, kind = 1, "avg"

    def __init__(self, pool_length=2, stride=None, border_mode="valid", input_shape=None, **kw):
        super().__init__((int(pool_length),), None if stride in (None, -1) else (int(stride),), border_mode, "tf", input_shape, **kw)


class MaxPooling2D(_Pool):
    nd, kind = 2, "max"

    def __init__(self, pool_size=(2, 2), strides=None, border_mode="valid", dim_ordering="th", input_shape=None, pads=None, **kw):
        super().__init__(pool_size, strides, border_mode, dim_ordering, input_shape, **kw)


class AveragePooling2D(_Pool):
    nd, kind = 2, "avg"

    def __init__(self, pool_size=(2, 2), strides=None, border_mode="valid", dim_ordering="th", input_shape=None, pads=None,
                 count_include_pad=False, **kw):
        super().__init__(pool_size, strides, border_mode, dim_ordering, input_shape, **kw)


class MaxPooling3D(_Pool):
    nd, kind = 3, "max"

    def __init__(self, pool_size=(2, 2, 2), strides=None, border_mode="valid", dim_ordering="th", input_shape=None, **kw):
        super().__init__(pool_size, strides, border_mode, dim_ordering, input_shape, **kw)


class AveragePooling3D(_Pool):
    nd, kind = 3, "avg"

    def __init__(self, pool_size=(2, 2, 2), strides=None, border_mode="valid", dim_ordering="th", input_shape=None, **kw):
        super().__init__(pool_size, strides, border_mode, dim_ordering, input_shape, **kw)


class _GlobalPool(_FnLayer):
    nd, kind = 2, "max"

    def __init__(self, dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dim_ordering = dim_ordering

    def fn(self, x):
        x = x.float()
        if self.nd == 1:
            dims = (1,)
        else:
            dims = tuple(range(2, 2 + self.nd)) if self.dim_ordering == "th" else tuple(range(1, 1 + self.nd))
        return x.amax(dims) if self.kind == "max" else x.mean(dims)


class GlobalMaxPooling1D(_GlobalPool):
    nd, kind = 1, "max"

    def __init__(self, input_shape=None, **kw):
        super().__init__("th", input_shape, **kw)


class GlobalAveragePooling1D(_GlobalPool):
    nd, kind = 1, "avg"

    def __init__(self, input_shape=None, **kw):
        super().__init__("th", input_shape, **kw)


class GlobalMaxPooling2D(_GlobalPool):
    nd, kind = 2, "max"


class GlobalAveragePooling2D(_GlobalPool):
    nd, kind = 2, "avg"


class GlobalMaxPooling3D(_GlobalPool):
    nd, kind = 3, "max"


class GlobalAveragePooling3D(_GlobalPool):
    nd, kind = 3, "avg"
