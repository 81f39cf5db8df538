"""Torch-style element / shape ops exposed as layers (reference: pyzoo/zoo/pipeline/api/keras/layers/torch.py)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .base import ZooKerasLayer, _FnLayer


def _dim(d, x):
    """Reference dims are 0-based over the FULL tensor (batch = 0); negative dims count from the end."""
    return d


class Select(_FnLayer):
    def __init__(self, dim, index, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dim, self.index = int(dim), int(index)

    def fn(self, x):
        return x.select(self.dim, self.index if self.index >= 0 else x.shape[self.dim] + self.index)


class Narrow(_FnLayer):
    def __init__(self, dim, offset, length=1, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dim, self.offset, self.length = int(dim), int(offset), int(length)

    def fn(self, x):
        n = x.shape[self.dim] - self.offset if self.length == -1 else self.length
        return x.narrow(self.dim, self.offset, n)


class Squeeze(_FnLayer):
    def __init__(self, dim=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dim = dim

    def fn(self, x):
        if self.dim is None:
            return x.squeeze()
        dims = self.dim if isinstance(self.dim, (list, tuple)) else [self.dim]
        for d in sorted(dims, reverse=True):
            x = x.squeeze(d)
        return x


class AddConstant(_FnLayer):
    def __init__(self, constant, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.constant = float(constant)

    def fn(self, x):
        return x + self.constant


class MulConstant(_FnLayer):
    def __init__(self, constant, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.constant = float(constant)

    def fn(self, x):
        return x * self.constant


class LRN2D(_FnLayer):
    def __init__(self, alpha=1e-4, k=1.0, beta=0.75, n=5, dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.args, self.dim_ordering = (int(n), float(alpha), float(beta), float(k)), dim_ordering

    def fn(self, x):
        if self.dim_ordering == "tf":
            return F.local_response_norm(x.permute(0, 3, 1, 2), *self.args).permute(0, 2, 3, 1)
        return F.local_response_norm(x, *self.args)


class WithinChannelLRN2D(_FnLayer):
    def __init__(self, size=5, alpha=1.0, beta=0.75, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size, self.alpha, self.beta = int(size), float(alpha), float(beta)

    def fn(self, x):
        sq = F.avg_pool2d(x * x, self.size, 1, self.size // 2, count_include_pad=True)
        return x / (1 + self.alpha * sq).pow(self.beta)


class CAdd(ZooKerasLayer):
    """Learnable bias of shape `size`, broadcast-added."""

    def __init__(self, size, b_regularizer=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size = tuple(int(s) for s in size)

    def build(self, input_shape):
        self.bias = nn.Parameter(torch.zeros(self.size))

    def call(self, x):
        return x + self.bias


class CMul(ZooKerasLayer):
    def __init__(self, size, W_regularizer=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size = tuple(int(s) for s in size)

    def build(self, input_shape):
        self.weight = nn.Parameter(torch.ones(self.size))

    def call(self, x):
        return x * self.weight


class Mul(ZooKerasLayer):
    """Multiply by one learnable scalar."""

    def build(self, input_shape):
        self.weight = nn.Parameter(torch.ones(1))

    def call(self, x):
        return x * self.weight


class Scale(ZooKerasLayer):
    def __init__(self, size, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size = tuple(int(s) for s in size)

    def build(self, input_shape):
        self.weight = nn.Parameter(torch.ones(self.size))
        self.bias = nn.Parameter(torch.zeros(self.size))

    def call(self, x):
        return x * self.weight + self.bias


class Exp(_FnLayer):
    def fn(self, x):
        return torch.exp(x)


class Identity(_FnLayer):
    def fn(self, x):
        return x


class Log(_FnLayer):
    def fn(self, x):
        return torch.log(x)


class Sqrt(_FnLayer):
    def fn(self, x):
        return torch.sqrt(x)


class Square(_FnLayer):
    def fn(self, x):
        return x * x


class Negative(_FnLayer):
    def fn(self, x):
        return -x


class Power(_FnLayer):
    """(shift + scale * x) ^ power."""

    def __init__(self, power, scale=1, shift=0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.power, self.scale, self.shift = float(power), float(scale), float(shift)

    def fn(self, x):
        return (self.shift + self.scale * x).pow(self.power)


class HardShrink(_FnLayer):
    def __init__(self, value=0.5, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.value = float(value)

    def fn(self, x):
        return F.hardshrink(x, self.value)


class SoftShrink(_FnLayer):
    def __init__(self, value=0.5, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.value = float(value)

    def fn(self, x):
        return F.softshrink(x, self.value)


class HardTanh(_FnLayer):
    def __init__(self, min_value=-1, max_value=1, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.lo, self.hi = float(min_value), float(max_value)

    def fn(self, x):
        return x.clamp(self.lo, self.hi)


class PReLU(ZooKerasLayer):
    def __init__(self, n_output_plane=0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.n = int(n_output_plane)

    def build(self, input_shape):
        self.act = nn.PReLU(max(self.n, 1))

    def call(self, x):
        return self.act(x)


class RReLU(_FnLayer):
    def __init__(self, lower=1.0 / 8, upper=1.0 / 3, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.lower, self.upper = float(lower), float(upper)

    def fn(self, x):
        return F.rrelu(x, self.lower, self.upper, self.training)


class BinaryThreshold(_FnLayer):
    def __init__(self, value=1e-6, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.value = float(value)

    def fn(self, x):
        return (x > self.value).to(x.dtype)


class Threshold(_FnLayer):
    def __init__(self, th=1e-6, v=0.0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.th, self.v = float(th), float(v)

    def fn(self, x):
        return torch.where(x > self.th, x, torch.full_like(x, self.v))


class GaussianSampler(_FnLayer):
    """Input [mean, log_variance] -> a sample of N(mean, exp(log_variance)) (VAE reparameterisation)."""
    multi_input = True

    def fn(self, xs):
        mean, logvar = xs
        return mean + torch.exp(0.5 * logvar) * torch.randn_like(mean)


class ResizeBilinear(_FnLayer):
    def __init__(self, output_height, output_width, align_corner=False, dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size, self.align, self.dim_ordering = (int(output_height), int(output_width)), bool(align_corner), dim_ordering

    def fn(self, x):
        if self.dim_ordering == "tf":
            x = x.permute(0, 3, 1, 2)
        y = F.interpolate(x.float(), size=self.size, mode="bilinear", align_corners=self.align)
        return y.permute(0, 2, 3, 1) if self.dim_ordering == "tf" else y


class SelectTable(_FnLayer):
    multi_input = True

    def __init__(self, index, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.index = int(index)

    def fn(self, xs):
        return xs[self.index]


class SplitTensor(_FnLayer):
    def __init__(self, dimension, num, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dimension, self.num = int(dimension), int(num)

    def fn(self, x):
        return list(x.chunk(self.num, self.dimension))


class Expand(_FnLayer):
    def __init__(self, tgt_sizes, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.tgt = tuple(int(s) for s in tgt_sizes)

    def fn(self, x):
        return x.expand(*[x.shape[0] if (i == 0 and s == -1) else s for i, s in enumerate(self.tgt)])


class Softmax(_FnLayer):
    def fn(self, x):
        return torch.softmax(x, -1)


class ShareConvolution2D(ZooKerasLayer):
    """Convolution2D whose weights can be shared (BigDL's memory-sharing variant; same math)."""

    def __init__(self, nb_filter, nb_row, nb_col, init="glorot_uniform", activation=None, subsample=(1, 1), pad_h=0, pad_w=0,
                 propagate_back=True, dim_ordering="th", W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.cfg = (int(nb_filter), (int(nb_row), int(nb_col)), tuple(subsample), (int(pad_h), int(pad_w)), bias)
        self.activation, self.init = activation, init

    def build(self, input_shape):
        from .base import get_activation, init_weight

        nf, k, s, p, b = self.cfg
        self.conv = nn.Conv2d(input_shape[0], nf, k, s, p, bias=b)
        init_weight(self.conv.weight, self.init)
        self._act = get_activation(self.activation)

    def call(self, x):
        return self._act(self.conv(x.float()))
