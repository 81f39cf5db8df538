"""Convolution / up-sampling / padding / cropping layers (reference: layers/convolutional.py; Scala Convolution1D/2D/3D,
AtrousConvolution1D/2D, Deconvolution2D, SeparableConvolution2D, UpSampling*, ZeroPadding*, Cropping*).
`dim_ordering="th"` (channels first, the reference default) or `"tf"`; 1-D layers take (batch, steps, dim)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .base import ZooKerasLayer, _FnLayer, get_activation, init_weight


def _pad_for(border_mode, k, dilation=1):
    if border_mode == "valid":
        return 0
    if border_mode == "same":
        return "same"
    raise ValueError(f"border_mode must be 'valid' or 'same', got {border_mode}")


class _ConvND(ZooKerasLayer):
    nd = 2

    def __init__(self, nb_filter, kernel, init, activation, border_mode, subsample, dilation, dim_ordering, bias, input_shape, **kw):
        super().__init__(input_shape, **kw)
        self.nb_filter, self.kernel, self.init, self.activation = int(nb_filter), kernel, init, activation
        self.border_mode, self.subsample, self.dilation = border_mode, subsample, dilation
        self.dim_ordering, self.use_bias = dim_ordering, bias

    def _cin(self, input_shape):
        return input_shape[0] if self.dim_ordering == "th" else input_shape[-1]

    def build(self, input_shape):
        cls = {1: nn.Conv1d, 2: nn.Conv2d, 3: nn.Conv3d}[self.nd]
        pad = _pad_for(self.border_mode, self.kernel)
        stride = self.subsample
        if pad == "same" and any(s != 1 for s in (stride if isinstance(stride, tuple) else (stride,))):
            pad = tuple((k - 1) * d // 2 for k, d in zip(self.kernel, self.dilation)) if isinstance(self.kernel, tuple) else (self.kernel - 1) // 2
        self.conv = cls(self._cin(input_shape), self.nb_filter, self.kernel, stride, pad, self.dilation, bias=self.use_bias)
        init_weight(self.conv.weight, self.init)
        self._act = get_activation(self.activation)

    def call(self, x):
        x = x.float()
        if self.dim_ordering == "tf":
            x = x.movedim(-1, 1)
        y = self._act(self.conv(x))
        return y.movedim(1, -1) if self.dim_ordering == "tf" else y


class Convolution1D(_ConvND):
    """Input (batch, steps, input_dim) -> (batch, new_steps, nb_filter)."""
    nd = 1

    def __init__(self, nb_filter, filter_length, init="glorot_uniform", activation=None, border_mode="valid", subsample_length=1,
                 W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(nb_filter, int(filter_length), init, activation, border_mode, int(subsample_length), 1, "tf", bias, input_shape, **kw)


class AtrousConvolution1D(_ConvND):
    nd = 1

    def __init__(self, nb_filter, filter_length, init="glorot_uniform", activation=None, border_mode="valid", subsample_length=1,
                 atrous_rate=1, W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        if border_mode != "valid":
            raise ValueError("For AtrousConvolution1D, only border_mode='valid' is supported for now")
        super().__init__(nb_filter, int(filter_length), init, activation, border_mode, int(subsample_length), int(atrous_rate), "tf", bias, input_shape, **kw)


class Convolution2D(_ConvND):
    nd = 2

    def __init__(self, nb_filter, nb_row, nb_col, init="glorot_uniform", activation=None, border_mode="valid", subsample=(1, 1),
                 dim_ordering="th", W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, pads=None, **kw):
        super().__init__(nb_filter, (int(nb_row), int(nb_col)), init, activation, border_mode, tuple(subsample), (1, 1), dim_ordering, bias, input_shape, **kw)


class AtrousConvolution2D(_ConvND):
    nd = 2

    def __init__(self, nb_filter, nb_row, nb_col, init="glorot_uniform", activation=None, border_mode="valid", subsample=(1, 1),
                 atrous_rate=(1, 1), dim_ordering="th", W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        if border_mode != "valid":
            raise ValueError("For AtrousConvolution2D, only border_mode='valid' is supported for now")
        super().__init__(nb_filter, (int(nb_row), int(nb_col)), init, activation, border_mode, tuple(subsample), tuple(atrous_rate), dim_ordering,
                         bias, input_shape, **kw)


class Convolution3D(_ConvND):
    nd = 3

    def __init__(self, nb_filter, kernel_dim1, kernel_dim2, kernel_dim3, init="glorot_uniform", activation=None, border_mode="valid",
                 subsample=(1, 1, 1), dim_ordering="th", W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(nb_filter, (int(kernel_dim1), int(kernel_dim2), int(kernel_dim3)), init, activation, border_mode, tuple(subsample),
                         (1, 1, 1), dim_ordering, bias, input_shape, **kw)


class Deconvolution2D(ZooKerasLayer):
    def __init__(self, nb_filter, nb_row, nb_col, output_shape=None, init="glorot_uniform", activation=None, border_mode="valid",
                 subsample=(1, 1), dim_ordering="th", W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if border_mode != "valid":
            raise ValueError("For Deconvolution2D, only border_mode='valid' is supported for now")
        self.cfg = (int(nb_filter), (int(nb_row), int(nb_col)), tuple(subsample), bias)
        self.init, self.activation = init, activation

    def build(self, input_shape):
        nf, k, s, b = self.cfg
        self.deconv = nn.ConvTranspose2d(input_shape[0], nf, k, s, bias=b)
        init_weight(self.deconv.weight, self.init)
        self._act = get_activation(self.activation)

    def call(self, x):
        return self._act(self.deconv(x.float()))


class SeparableConvolution2D(ZooKerasLayer):
    def __init__(self, nb_filter, nb_row, nb_col, init="glorot_uniform", activation=None, border_mode="valid", subsample=(1, 1),
                 depth_multiplier=1, dim_ordering="th", depthwise_regularizer=None, pointwise_regularizer=None, b_regularizer=None,
                 bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.cfg = (int(nb_filter), (int(nb_row), int(nb_col)), tuple(subsample), int(depth_multiplier), bias, border_mode)
        self.init, self.activation, self.dim_ordering = init, activation, dim_ordering

    def build(self, input_shape):
        nf, k, s, dm, b, bm = self.cfg
        cin = input_shape[0] if self.dim_ordering == "th" else input_shape[-1]
        pad = 0 if bm == "valid" else tuple((kk - 1) // 2 for kk in k)
        self.depthwise = nn.Conv2d(cin, cin * dm, k, s, pad, groups=cin, bias=False)
        self.pointwise = nn.Conv2d(cin * dm, nf, 1, bias=b)
        init_weight(self.depthwise.weight, self.init)
        init_weight(self.pointwise.weight, self.init)
        self._act = get_activation(self.activation)

    def call(self, x):
        x = x.float()
        if self.dim_ordering == "tf":
            x = x.movedim(-1, 1)
        y = self._act(self.pointwise(self.depthwise(x)))
        return y.movedim(1, -1) if self.dim_ordering == "tf" else y


class UpSampling1D(_FnLayer):
    def __init__(self, length=2, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.length = int(length)

    def fn(self, x):
        return x.repeat_interleave(self.length, dim=1)


class UpSampling2D(_FnLayer):
    def __init__(self, size=(2, 2), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size, self.dim_ordering = tuple(size), dim_ordering

    def fn(self, x):
        a, b = (2, 3) if self.dim_ordering == "th" else (1, 2)
        return x.repeat_interleave(self.size[0], dim=a).repeat_interleave(self.size[1], dim=b)


class UpSampling3D(_FnLayer):
    def __init__(self, size=(2, 2, 2), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size, self.dim_ordering = tuple(size), dim_ordering

    def fn(self, x):
        dims = (2, 3, 4) if self.dim_ordering == "th" else (1, 2, 3)
        for d, s in zip(dims, self.size):
            x = x.repeat_interleave(s, dim=d)
        return x


class ZeroPadding1D(_FnLayer):
    def __init__(self, padding=1, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.padding = (padding, padding) if isinstance(padding, int) else tuple(padding)

    def fn(self, x):
        return F.pad(x, (0, 0, self.padding[0], self.padding[1]))


class ZeroPadding2D(_FnLayer):
    def __init__(self, padding=(1, 1), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        p = tuple(padding)
        self.padding = (p[0], p[0], p[1], p[1]) if len(p) == 2 else p     # top, bottom, left, right
        self.dim_ordering = dim_ordering

    def fn(self, x):
        t, b, l, r = self.padding
        return F.pad(x, (l, r, t, b)) if self.dim_ordering == "th" else F.pad(x, (0, 0, l, r, t, b))


class ZeroPadding3D(_FnLayer):
    def __init__(self, padding=(1, 1, 1), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.padding, self.dim_ordering = tuple(padding), dim_ordering

    def fn(self, x):
        a, b, c = self.padding
        return F.pad(x, (c, c, b, b, a, a)) if self.dim_ordering == "th" else F.pad(x, (0, 0, c, c, b, b, a, a))


class Cropping1D(_FnLayer):
    def __init__(self, cropping=(1, 1), input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.cropping = tuple(cropping)

    def fn(self, x):
        return x[:, self.cropping[0]:x.shape[1] - self.cropping[1]]


class Cropping2D(_FnLayer):
    def __init__(self, cropping=((0, 0), (0, 0)), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.cropping, self.dim_ordering = tuple(tuple(c) for c in cropping), dim_ordering

    def fn(self, x):
        (t, b), (l, r) = self.cropping
        if self.dim_ordering == "th":
            return x[:, :, t:x.shape[2] - b, l:x.shape[3] - r]
        return x[:, t:x.shape[1] - b, l:x.shape[2] - r]


class Cropping3D(_FnLayer):
    def __init__(self, cropping=((1, 1), (1, 1), (1, 1)), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.cropping, self.dim_ordering = tuple(tuple(c) for c in cropping), dim_ordering

    def fn(self, x):
        (a0, a1), (b0, b1), (c0, c1) = self.cropping
        if self.dim_ordering == "th":
            return x[:, :, a0:x.shape[2] - a1, b0:x.shape[3] - b1, c0:x.shape[4] - c1]
        return x[:, a0:x.shape[1] - a1, b0:x.shape[2] - b1, c0:x.shape[3] - c1]
