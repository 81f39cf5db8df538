"""BatchNormalization / LayerNorm (reference: layers/normalization.py, Scala BatchNormalization.scala, LayerNorm.scala)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .base import ZooKerasLayer


class BatchNormalization(ZooKerasLayer):
    """Per-channel batch norm: 4-D input normalises over (batch, H, W) (`dim_ordering` picks the channel axis),
    2-D input over the batch.  `momentum` is Keras' (running = momentum*running + (1-momentum)*batch)."""

    def __init__(self, epsilon=1e-3, mode=0, axis=1, momentum=0.99, beta_init="zero", gamma_init="one", dim_ordering="th",
                 input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.epsilon, self.momentum, self.dim_ordering = float(epsilon), float(momentum), dim_ordering

    def build(self, input_shape):
        nd = len(input_shape)
        c = input_shape[0] if (self.dim_ordering == "th" or nd == 1) else input_shape[-1]
        cls = nn.BatchNorm1d if nd <= 2 else (nn.BatchNorm2d if nd == 3 else nn.BatchNorm3d)
        self.bn = cls(c, eps=self.epsilon, momentum=1.0 - self.momentum)
        self.nd = nd

    def call(self, x):
        x = x.float()
        if self.dim_ordering == "tf" and self.nd > 1:
            return self.bn(x.movedim(-1, 1)).movedim(1, -1)
        return self.bn(x)


class LayerNorm(ZooKerasLayer):
    def __init__(self, n_output=None, e_value=1e-5, hidden_size=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.n, self.eps = n_output or hidden_size, float(e_value)

    def build(self, input_shape):
        self.ln = nn.LayerNorm(self.n or input_shape[-1], eps=self.eps)

    def call(self, x):
        return self.ln(x.float())
