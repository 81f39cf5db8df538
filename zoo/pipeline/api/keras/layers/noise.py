"""GaussianNoise / GaussianDropout (reference: layers/noise.py)."""
from __future__ import annotations

import math

import torch

from .base import _FnLayer


class GaussianNoise(_FnLayer):
    def __init__(self, sigma, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.sigma = float(sigma)

    def fn(self, x):
        return x + torch.randn_like(x) * self.sigma if self.training else x


class GaussianDropout(_FnLayer):
    def __init__(self, p, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.p = float(p)

    def fn(self, x):
        if not self.training:
            return x
        return x * (1.0 + torch.randn_like(x) * math.sqrt(self.p / (1.0 - self.p)))
