"""Validation methods.

Reference: zoo/src/main/scala/.../pipeline/api/keras/metrics/{Accuracy,AUC,MAE}.scala (366 LoC)
plus the BigDL methods the reference re-exports (Top1Accuracy, Top5Accuracy, Loss).  A metric
accumulates `(numerator, denominator)` style statistics on the device batch by batch; ranks are
merged by summation (``state()`` / ``merge()``), which is how distributed `evaluate` reduces.
"""
from __future__ import annotations

import numpy as np
import torch


class ValidationResult:
    def __init__(self, name, value, count):
        self.name, self.value, self.count = name, float(value), int(count)

    def result(self):
        return self.value, self.count

    def __repr__(self):
        return f"{self.name}: {self.value:.6f} (count {self.count})"


class ValidationMethod:
    name = "metric"

    def __init__(self):
        self.reset()

    def reset(self):
        self.acc = torch.zeros(2, dtype=torch.float64)

    def update(self, output, target, loss=None):
        raise NotImplementedError

    def _add(self, num, den):
        self.acc += torch.tensor([float(num), float(den)], dtype=torch.float64)

    def state(self):
        return self.acc.clone()

    def load(self, state):
        self.acc = state.clone().double().cpu()

    def result(self):
        n, d = self.acc.tolist()
        return ValidationResult(self.name, n / d if d else 0.0, d)

    def __repr__(self):
        return self.name


def _labels(target, zero_based):
    t = target.reshape(-1).long()
    return t if zero_based else t - 1


class Accuracy(ValidationMethod):
    """Top-1 accuracy; binary when the output has one column (threshold 0.5), sparse integer labels
    or one-hot targets otherwise (metrics/Accuracy.scala)."""
    name = "Top1Accuracy"

    def __init__(self, zero_based_label=True):
        self.zero_based_label = zero_based_label
        super().__init__()

    def update(self, output, target, loss=None):
        o = output.detach().float()
        o = o.reshape(-1, o.shape[-1]) if o.dim() > 1 else o.reshape(-1, 1)
        if o.shape[-1] == 1:
            pred = (o.reshape(-1) > 0.5).long()
            t = target.reshape(-1).long()
        else:
            pred = o.argmax(-1)
            t = target.argmax(-1).reshape(-1) if (target.dim() > 1 and target.shape[-1] == o.shape[-1] and target.shape[-1] > 1) \
                else _labels(target, self.zero_based_label)
        self._add((pred == t.to(pred.device)).sum().item(), pred.numel())


Top1Accuracy = SparseCategoricalAccuracy = CategoricalAccuracy = BinaryAccuracy = Accuracy


class Top5Accuracy(ValidationMethod):
    name = "Top5Accuracy"

    def __init__(self, zero_based_label=True):
        self.zero_based_label = zero_based_label
        super().__init__()

    def update(self, output, target, loss=None):
        o = output.detach().float().reshape(-1, output.shape[-1])
        k = min(5, o.shape[-1])
        top = o.topk(k, dim=-1).indices
        t = _labels(target, self.zero_based_label).to(top.device)
        self._add((top == t.unsqueeze(-1)).any(-1).sum().item(), t.numel())


class MAE(ValidationMethod):
    name = "MAE"

    def update(self, output, target, loss=None):
        o = output.detach().float()
        self._add((o - target.float().reshape(o.shape).to(o.device)).abs().sum().item(), o.numel())


class MSE(ValidationMethod):
    name = "MSE"

    def update(self, output, target, loss=None):
        o = output.detach().float()
        self._add(((o - target.float().reshape(o.shape).to(o.device)) ** 2).sum().item(), o.numel())


class Loss(ValidationMethod):
    """Mean of the criterion over validation batches (BigDL `Loss`)."""
    name = "Loss"

    def __init__(self, criterion=None):
        self.criterion = criterion
        super().__init__()

    def update(self, output, target, loss=None):
        if loss is None and self.criterion is not None:
            loss = self.criterion(output, target)
        n = output.shape[0]
        self._add(float(loss) * n, n)


class AUC(ValidationMethod):
    """Area under the ROC curve with `threshold_num` buckets (metrics/AUC.scala)."""
    name = "AUC"

    def __init__(self, threshold_num=200):
        self.threshold_num = int(threshold_num)
        super().__init__()

    def reset(self):
        self.acc = torch.zeros(2, self.threshold_num, dtype=torch.float64)

    def update(self, output, target, loss=None):
        o = output.detach().float()
        p = o[:, -1] if (o.dim() > 1 and o.shape[-1] == 2) else o.reshape(-1)
        t = target.reshape(-1).float().to(p.device)
        b = (p.clamp(0, 1) * (self.threshold_num - 1)).long().cpu()
        tc = t.cpu().double()
        self.acc[0].index_add_(0, b, tc)
        self.acc[1].index_add_(0, b, 1.0 - tc)

    def result(self):
        pos, neg = self.acc[0].flip(0).cumsum(0), self.acc[1].flip(0).cumsum(0)
        P, N = pos[-1].item(), neg[-1].item()
        if P == 0 or N == 0:
            return ValidationResult(self.name, 0.0, P + N)
        tpr = torch.cat([torch.zeros(1, dtype=torch.float64), pos / P])
        fpr = torch.cat([torch.zeros(1, dtype=torch.float64), neg / N])
        return ValidationResult(self.name, torch.trapz(tpr, fpr).item(), P + N)


class TorchMetric(ValidationMethod):
    """Wrap `fn(output, target) -> scalar` averaged over samples."""

    def __init__(self, fn, name=None):
        self.fn, self.name = fn, name or getattr(fn, "__name__", "metric")
        super().__init__()

    def update(self, output, target, loss=None):
        n = output.shape[0]
        self._add(float(self.fn(output, target)) * n, n)


def to_metric(m, loss=None):
    if isinstance(m, ValidationMethod):
        return m
    if isinstance(m, str):
        k = m.lower()
        if k in ("accuracy", "acc", "top1accuracy", "top1"):
            zb = getattr(loss, "zero_based_label", True)
            return Accuracy(zero_based_label=zb)
        if k in ("top5accuracy", "top5acc", "top5"):
            return Top5Accuracy(zero_based_label=getattr(loss, "zero_based_label", True))
        if k == "mae":
            return MAE()
        if k == "mse":
            return MSE()
        if k == "auc":
            return AUC()
        if k == "loss":
            return Loss(loss)
        raise ValueError(f"unsupported metric {m!r}")
    if callable(m):
        return TorchMetric(m)
    raise TypeError(f"cannot interpret {m!r} as a metric")
