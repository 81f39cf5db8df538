"""L1 / L2 weight regularizers (reference: pyzoo/zoo/pipeline/api/keras/regularizers.py -> BigDL L1L2Regularizer).
A layer given `W_regularizer=l2(1e-4)` adds `l2 * sum(w^2) + l1 * sum(|w|)` of that weight to the training loss."""
from __future__ import annotations


class L1L2Regularizer:
    def __init__(self, l1=0.0, l2=0.0):
        self.l1, self.l2 = float(l1), float(l2)

    def __call__(self, w):
        out = 0.0
        if self.l1:
            out = out + self.l1 * w.abs().sum()
        if self.l2:
            out = out + self.l2 * (w * w).sum()
        return out

    def __repr__(self):
        return f"L1L2Regularizer(l1={self.l1}, l2={self.l2})"


def l1(l1=0.01):
    return L1L2Regularizer(l1=l1)


def l2(l2=0.01):
    return L1L2Regularizer(l2=l2)


def l1l2(l1=0.01, l2=0.01):
    return L1L2Regularizer(l1=l1, l2=l2)
