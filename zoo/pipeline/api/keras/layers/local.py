"""LocallyConnected1D / 2D: convolution with unshared weights (reference: layers/local.py)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .base import ZooKerasLayer, get_activation, init_weight


class LocallyConnected1D(ZooKerasLayer):
    """(batch, steps, dim) -> (batch, new_steps, nb_filter); border_mode 'valid' only."""

    def __init__(self, nb_filter, filter_length, activation=None, border_mode="valid", subsample_length=1, W_regularizer=None,
                 b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if border_mode != "valid":
            raise ValueError("For LocallyConnected1D, only border_mode='valid' is supported for now")
        self.nb_filter, self.k, self.s, self.activation, self.use_bias = int(nb_filter), int(filter_length), int(subsample_length), activation, bias

    def build(self, input_shape):
        steps, dim = input_shape
        self.out_steps = (steps - self.k) // self.s + 1
        self.weight = nn.Parameter(init_weight(torch.empty(self.out_steps, self.k * dim, self.nb_filter)))
        self.bias = nn.Parameter(torch.zeros(self.out_steps, self.nb_filter)) if self.use_bias else None
        self._act = get_activation(self.activation)

    def call(self, x):
        p = x.float().unfold(1, self.k, self.s)                    # B, out_steps, dim, k
        p = p.permute(0, 1, 3, 2).reshape(x.shape[0], self.out_steps, -1)
        y = torch.einsum("bsk,skf->bsf", p, self.weight)
        return self._act(y + self.bias if self.bias is not None else y)


class LocallyConnected2D(ZooKerasLayer):
    def __init__(self, nb_filter, nb_row, nb_col, activation=None, border_mode="valid", subsample=(1, 1), dim_ordering="th",
                 W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.nb_filter, self.k, self.s = int(nb_filter), (int(nb_row), int(nb_col)), tuple(subsample)
        self.activation, self.border_mode, self.dim_ordering, self.use_bias = activation, border_mode, dim_ordering, bias

    def _pad(self):
        return (0, 0) if self.border_mode == "valid" else ((self.k[0] - 1) // 2, (self.k[1] - 1) // 2)

    def build(self, input_shape):
        c, h, w = input_shape if self.dim_ordering == "th" else (input_shape[2], input_shape[0], input_shape[1])
        ph, pw = self._pad()
        self.oh = (h + 2 * ph - self.k[0]) // self.s[0] + 1
        self.ow = (w + 2 * pw - self.k[1]) // self.s[1] + 1
        self.weight = nn.Parameter(init_weight(torch.empty(self.oh * self.ow, c * self.k[0] * self.k[1], self.nb_filter)))
        self.bias = nn.Parameter(torch.zeros(self.oh * self.ow, self.nb_filter)) if self.use_bias else None
        self._act = get_activation(self.activation)

    def call(self, x):
        x = x.float()
        if self.dim_ordering == "tf":
            x = x.permute(0, 3, 1, 2)
        p = F.unfold(x, self.k, padding=self._pad(), stride=self.s).transpose(1, 2)      # B, L, C*kh*kw
        y = torch.einsum("blk,lkf->blf", p, self.weight)
        if self.bias is not None:
            y = y + self.bias
        y = self._act(y).reshape(x.shape[0], self.oh, self.ow, self.nb_filter)
        return y if self.dim_ordering == "tf" else y.permute(0, 3, 1, 2)
