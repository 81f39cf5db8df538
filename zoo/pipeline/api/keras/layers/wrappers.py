"""TimeDistributed / Bidirectional / KerasLayerWrapper (reference: layers/wrappers.py)."""
from __future__ import annotations

import copy

import torch
import torch.nn as nn

from .base import ZooKerasLayer


class TimeDistributed(ZooKerasLayer):
    """Apply `layer` to every timestep of (batch, steps, ...)."""

    def __init__(self, layer, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.layer = layer

    def build(self, input_shape):
        if isinstance(self.layer, ZooKerasLayer):
            self.layer.ensure_built(tuple(input_shape[1:]))

    def call(self, x):
        B, T = x.shape[:2]
        y = self.layer(x.reshape(B * T, *x.shape[2:]))
        return y.reshape(B, T, *y.shape[1:])


class Bidirectional(ZooKerasLayer):
    """Run a recurrent layer forwards and backwards and merge: concat | sum | mul | ave."""

    def __init__(self, layer, merge_mode="concat", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if merge_mode not in ("concat", "sum", "mul", "ave"):
            raise ValueError(f"Invalid merge mode: {merge_mode}")
        self.fwd, self.merge_mode = layer, merge_mode
        self.bwd = copy.deepcopy(layer)
        self.bwd.go_backwards = not getattr(layer, "go_backwards", False)
        self.bwd.zoo_name = self.fwd.zoo_name + "_reverse"

    def build(self, input_shape):
        self.fwd.ensure_built(input_shape)
        self.bwd.ensure_built(input_shape)

    def call(self, x):
        a, b = self.fwd(x), self.bwd(x)
        if getattr(self.fwd, "return_sequences", False):
            b = b.flip(1)
        if self.merge_mode == "concat":
            return torch.cat([a, b], -1)
        if self.merge_mode == "sum":
            return a + b
        if self.merge_mode == "mul":
            return a * b
        return (a + b) / 2


class KerasLayerWrapper(ZooKerasLayer):
    """Wrap any torch module as a zoo keras layer (the reference wraps a BigDL layer)."""

    def __init__(self, torch_layer: nn.Module, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.layer = torch_layer

    def call(self, x):
        return self.layer(*x) if isinstance(x, list) else self.layer(x)
