"""LeakyReLU / ELU / ThresholdedReLU / SReLU (reference: layers/advanced_activations.py)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .base import ZooKerasLayer, _FnLayer


class LeakyReLU(_FnLayer):
    def __init__(self, alpha=0.01, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.alpha = float(alpha)

    def fn(self, x):
        return F.leaky_relu(x, self.alpha)


class ELU(_FnLayer):
    def __init__(self, alpha=1.0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.alpha = float(alpha)

    def fn(self, x):
        return F.elu(x, self.alpha)


class ThresholdedReLU(_FnLayer):
    def __init__(self, theta=1.0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.theta = float(theta)

    def fn(self, x):
        return x * (x > self.theta).to(x.dtype)


class SReLU(ZooKerasLayer):
    """S-shaped ReLU with learnable (t_left, a_left, t_right, a_right) per feature."""

    def __init__(self, t_left_init="zero", a_left_init="glorot_uniform", t_right_init="glorot_uniform", a_right_init="one",
                 shared_axes=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.inits, self.shared_axes = (t_left_init, a_left_init, t_right_init, a_right_init), shared_axes

    def build(self, input_shape):
        from .base import init_weight

        shape = list(input_shape)
        for ax in (self.shared_axes or []):
            shape[ax - 1] = 1
        self.t_left, self.a_left, self.t_right, self.a_right = [nn.Parameter(init_weight(torch.empty(shape), i)) for i in self.inits]

    def call(self, x):
        tr = self.t_left + self.t_right.abs()
        y = torch.where(x >= tr, tr + self.a_right * (x - tr), x)
        return torch.where(x <= self.t_left, self.t_left + self.a_left * (x - self.t_left), y)
