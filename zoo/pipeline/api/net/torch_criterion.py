"""TorchCriterion: a PyTorch loss as a zoo criterion (reference: pyzoo/zoo/pipeline/api/net/torch_criterion.py:30-93,
TorchCriterion.scala:30-129 -- a traced `LossWrapperModule` run through JNI).  Here the callable is used directly."""
from __future__ import annotations

import torch
import torch.nn as nn

from ..keras.objectives import LossFunction


class LossWrapperModule(nn.Module):
    def __init__(self, lossFunc):
        super().__init__()
        self.func = lossFunc

    def forward(self, x, y):
        return self.func(x, y)


class TorchCriterion(LossFunction):
    def __init__(self, loss_or_path, bigdl_type="float"):
        super().__init__()
        self.func = loss_or_path if callable(loss_or_path) else torch.jit.load(loss_or_path, map_location="cpu")

    @staticmethod
    def from_pytorch(loss, input=None, label=None):
        """`loss`: an nn loss module or `f(input, target)`.  `input` / `label` (sample shapes) are accepted for
        API parity; they were only needed for tracing."""
        return TorchCriterion(loss)

    def forward(self, y_pred, y_true):
        t = y_true
        if isinstance(self.func, (nn.CrossEntropyLoss, nn.NLLLoss)) and torch.is_tensor(t):
            t = t.reshape(-1).long() if t.dim() > 1 and t.shape[-1] == 1 else t.long() if not t.is_floating_point() or t.dim() == 1 else t
        elif torch.is_tensor(t) and torch.is_tensor(y_pred) and t.is_floating_point() is False and y_pred.is_floating_point() \
                and not isinstance(self.func, (nn.CrossEntropyLoss, nn.NLLLoss)) and isinstance(self.func, nn.Module):
            t = t.to(y_pred.dtype)
        out = self.func(y_pred.float() if torch.is_tensor(y_pred) else y_pred, t)
        return out.mean() if out.dim() else out
