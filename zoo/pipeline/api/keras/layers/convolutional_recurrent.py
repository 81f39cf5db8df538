"""ConvLSTM2D / ConvLSTM3D (reference: layers/convolutional_recurrent.py; Scala ConvLSTM2D.scala, InternalConvLSTM2D.scala).
Input (batch, time, channels, H, W[, D]) channels-first; square kernels; border_mode 'same'."""
from __future__ import annotations

import torch
import torch.nn as nn

from .base import ZooKerasLayer, get_activation


class _ConvLSTM(ZooKerasLayer):
    nd = 2

    def __init__(self, nb_filter, nb_kernel, activation="tanh", inner_activation="hard_sigmoid", dim_ordering="th", border_mode="same",
                 subsample=1, W_regularizer=None, U_regularizer=None, b_regularizer=None, return_sequences=False, go_backwards=False,
                 input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if dim_ordering != "th":
            raise ValueError("ConvLSTM only supports dim_ordering='th' (channels first)")
        if border_mode != "same":
            raise ValueError("ConvLSTM only supports border_mode='same'")
        self.nb_filter, self.k = int(nb_filter), int(nb_kernel)
        self.subsample = subsample[0] if isinstance(subsample, (tuple, list)) else int(subsample)
        self.act, self.inner = get_activation(activation), get_activation(inner_activation)
        self.return_sequences, self.go_backwards = return_sequences, go_backwards

    def build(self, input_shape):
        cin = input_shape[1]
        conv = nn.Conv2d if self.nd == 2 else nn.Conv3d
        self.wx = conv(cin, 4 * self.nb_filter, self.k, self.subsample, self.k // 2)
        self.wh = conv(self.nb_filter, 4 * self.nb_filter, self.k, 1, self.k // 2, bias=False)

    def call(self, x):
        x = x.float()
        if self.go_backwards:
            x = x.flip(1)
        h = c = None
        outs = []
        for t in range(x.shape[1]):
            g = self.wx(x[:, t])
            if h is not None:
                g = g + self.wh(h)
            i, f, gg, o = g.chunk(4, 1)
            cn = self.inner(i) * self.act(gg)
            c = cn if c is None else self.inner(f) * c + cn
            h = self.inner(o) * self.act(c)
            outs.append(h)
        return torch.stack(outs, 1) if self.return_sequences else h


class ConvLSTM2D(_ConvLSTM):
    nd = 2


class ConvLSTM3D(_ConvLSTM):
    nd = 3
