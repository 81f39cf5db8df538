"""Loss functions (the 14 objectives of the reference keras API).

Reference: zoo/src/main/scala/.../pipeline/api/keras/objectives/*.scala (940 LoC) and
pyzoo/zoo/pipeline/api/keras/objectives.py.  Every objective is an ``nn.Module`` taking
``(y_pred, y_true)``.  `SparseCategoricalCrossEntropy` and `BinaryCrossEntropy` additionally
implement ``from_logits(logits, target)`` which runs the fused sm_100a loss kernels
(``zoo.ops.loss``); the training engine calls that when the model exposes
``forward_logits`` so SoftMax + clip + log + NLL + gradient is one kernel launch.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from ....ops import loss as _fl

EPS = 1e-8


class LossFunction(nn.Module):
    accepts_logits = False
    logits_kind = None  # "softmax" | "sigmoid": which final activation `from_logits` absorbs

    def __init__(self, size_average=True):
        super().__init__()
        self.size_average = size_average

    def _reduce(self, per):
        return per.mean() if self.size_average else per.sum()


class SparseCategoricalCrossEntropy(LossFunction):
    """y_pred: probabilities [B, C] (or log-probs with log_prob_as_input); y_true: integer labels.
    SparseCategoricalCrossEntropy.scala:47-80 -> ZooClassNLLCriterion.scala:48-121."""
    accepts_logits = True
    logits_kind = "softmax"

    def __init__(self, log_prob_as_input=False, zero_based_label=True, weights=None, size_average=True, padding_value=-1):
        super().__init__(size_average)
        self.log_prob_as_input, self.zero_based_label = log_prob_as_input, zero_based_label
        self.padding_value = padding_value
        self.register_buffer("weights", None if weights is None else torch.as_tensor(weights, dtype=torch.float32))

    def forward(self, y_pred, y_true):
        t = y_true.reshape(-1).long() - (0 if self.zero_based_label else 1)
        p = y_pred.reshape(-1, y_pred.shape[-1]).float()
        valid = (y_true.reshape(-1).long() != self.padding_value) if self.padding_value is not None else torch.ones_like(t, dtype=torch.bool)
        tc = t.clamp(0, p.shape[-1] - 1)
        picked = p.gather(-1, tc.unsqueeze(-1)).squeeze(-1)
        logp = picked if self.log_prob_as_input else torch.log(picked.clamp(EPS, 1.0 - EPS))
        w = self.weights[tc] if self.weights is not None else torch.ones_like(logp)
        per = torch.where(valid, -w * logp, torch.zeros_like(logp))
        if self.size_average:
            return per.sum() / valid.sum().clamp(min=1)
        return per.sum()

    def from_logits(self, logits, y_true):
        return _fl.softmax_nll(logits.reshape(-1, logits.shape[-1]), y_true, self.zero_based_label, self.padding_value,
                               self.weights, self.size_average)


class CategoricalCrossEntropy(LossFunction):
    """y_true one-hot (CategoricalCrossEntropy.scala): -sum(t * log(clip(p)))."""
    accepts_logits = True
    logits_kind = "softmax"

    def forward(self, y_pred, y_true):
        p = y_pred.float().clamp(EPS, 1.0 - EPS)
        return self._reduce(-(y_true.float() * torch.log(p)).sum(-1))

    def from_logits(self, logits, y_true):
        return _fl.softmax_nll(logits.reshape(-1, logits.shape[-1]), y_true.argmax(-1), True, None, None, self.size_average)


class BinaryCrossEntropy(LossFunction):
    """BinaryCrossEntropy.scala:36-42 (BCECriterion on probabilities)."""
    accepts_logits = True
    logits_kind = "sigmoid"

    def __init__(self, weights=None, size_average=True):
        super().__init__(size_average)
        self.register_buffer("weights", None if weights is None else torch.as_tensor(weights, dtype=torch.float32))

    def forward(self, y_pred, y_true):
        if self.weights is None:
            return _fl.sigmoid_bce(y_pred.reshape(-1), y_true.reshape(-1), from_probs=True, size_average=self.size_average)
        p = y_pred.float()
        y = y_true.float()
        return self._reduce(-self.weights * (y * torch.log(p.clamp_min(EPS)) + (1 - y) * torch.log((1 - p).clamp_min(EPS))))

    def from_logits(self, logits, y_true):
        return _fl.sigmoid_bce(logits.reshape(-1), y_true.reshape(-1), from_probs=False, size_average=self.size_average)


class MeanSquaredError(LossFunction):
    def forward(self, y_pred, y_true):
        return self._reduce((y_pred.float() - y_true.float().reshape(y_pred.shape)) ** 2)


class MeanAbsoluteError(LossFunction):
    def forward(self, y_pred, y_true):
        return self._reduce((y_pred.float() - y_true.float().reshape(y_pred.shape)).abs())


class MeanAbsolutePercentageError(LossFunction):
    def forward(self, y_pred, y_true):
        y = y_true.float().reshape(y_pred.shape)
        return 100.0 * self._reduce(((y - y_pred.float()) / y.abs().clamp(EPS, None)).abs())


class MeanSquaredLogarithmicError(LossFunction):
    def forward(self, y_pred, y_true):
        y = y_true.float().reshape(y_pred.shape)
        a = torch.log(y_pred.float().clamp(EPS, None) + 1.0)
        b = torch.log(y.clamp(EPS, None) + 1.0)
        return self._reduce((a - b) ** 2)


class Hinge(LossFunction):
    def __init__(self, margin=1.0, size_average=True):
        super().__init__(size_average)
        self.margin = margin

    def forward(self, y_pred, y_true):
        y = y_true.float().reshape(y_pred.shape)
        return self._reduce(torch.clamp(self.margin - y * y_pred.float(), min=0))


class SquaredHinge(Hinge):
    def forward(self, y_pred, y_true):
        y = y_true.float().reshape(y_pred.shape)
        return self._reduce(torch.clamp(self.margin - y * y_pred.float(), min=0) ** 2)


class RankHinge(LossFunction):
    """Pairwise ranking hinge for text matching (RankHinge.scala): the batch interleaves
    (positive, negative) rows; loss = max(0, margin - pos + neg)."""

    def __init__(self, margin=1.0, size_average=True):
        super().__init__(size_average)
        self.margin = margin

    def forward(self, y_pred, y_true=None):
        p = y_pred.float().reshape(-1)
        pos, neg = p[0::2], p[1::2]
        return self._reduce(torch.clamp(self.margin - pos + neg, min=0))


class KullbackLeiblerDivergence(LossFunction):
    def forward(self, y_pred, y_true):
        p = y_pred.float().clamp(EPS, 1.0)
        t = y_true.float().clamp(EPS, 1.0)
        return self._reduce((t * torch.log(t / p)).sum(-1))


class Poisson(LossFunction):
    def forward(self, y_pred, y_true):
        p = y_pred.float()
        return self._reduce(p - y_true.float().reshape(p.shape) * torch.log(p + EPS))


class CosineProximity(LossFunction):
    def forward(self, y_pred, y_true):
        a = F.normalize(y_pred.float(), dim=-1)
        b = F.normalize(y_true.float().reshape(y_pred.shape), dim=-1)
        return self._reduce(-(a * b).sum(-1))


class TorchLoss(LossFunction):
    """Wrap any torch criterion / callable (the role of TorchCriterion.from_pytorch)."""

    def __init__(self, fn):
        super().__init__()
        self.fn = fn
        # a Keras / zoo style function names its first argument y_true (reference: compile wraps such a function in CustomLoss, which
        # calls it as (y_true, y_pred)); torch criteria take (input, target)
        try:
            import inspect

            first = next(iter(inspect.signature(fn).parameters), "")
        except (TypeError, ValueError):
            first = ""
        self.true_first = first in ("y_true", "ytrue", "target", "targets", "labels", "label")

    def forward(self, y_pred, y_true):
        out = self.fn(y_true, y_pred) if self.true_first else self.fn(y_pred, y_true)
        return out.mean() if torch.is_tensor(out) and out.dim() > 0 else out      # per-sample losses are averaged over the batch


_ALIASES = {
    "sparse_categorical_crossentropy": SparseCategoricalCrossEntropy, "categorical_crossentropy": CategoricalCrossEntropy,
    "binary_crossentropy": BinaryCrossEntropy, "mse": MeanSquaredError, "mean_squared_error": MeanSquaredError,
    "mae": MeanAbsoluteError, "mean_absolute_error": MeanAbsoluteError, "mape": MeanAbsolutePercentageError,
    "mean_absolute_percentage_error": MeanAbsolutePercentageError, "msle": MeanSquaredLogarithmicError,
    "mean_squared_logarithmic_error": MeanSquaredLogarithmicError, "hinge": Hinge, "squared_hinge": SquaredHinge,
    "rank_hinge": RankHinge, "kld": KullbackLeiblerDivergence, "kullback_leibler_divergence": KullbackLeiblerDivergence,
    "poisson": Poisson, "cosine_proximity": CosineProximity,
}


def to_loss(loss):
    """str | LossFunction | torch criterion | callable -> nn.Module (KerasUtils.toBigDLCriterion)."""
    if isinstance(loss, LossFunction):
        return loss
    if isinstance(loss, str):
        try:
            return _ALIASES[loss.lower()]()
        except KeyError:
            raise ValueError(f"unsupported loss {loss!r}") from None
    if callable(loss):
        return TorchLoss(loss)
    raise TypeError(f"cannot interpret {loss!r} as a loss")
