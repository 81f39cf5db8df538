"""Ranking metrics for text-matching style models.

Reference: zoo/src/main/scala/.../models/common/Ranker.scala:33-175 (`evaluateNDCG`,
`evaluateMAP` over a TextSet whose every sample is one query's batch of (features, labels))
and pyzoo/zoo/models/common/ranker.py.
"""
from __future__ import annotations

import numpy as np
import torch


def ndcg(y_true, y_pred, k, threshold=0.0):
    """NDCG@k of one query (Ranker.scala:120-150): gain 2^label - 1, log2 discount."""
    y_true, y_pred = np.asarray(y_true, dtype=np.float64).ravel(), np.asarray(y_pred, dtype=np.float64).ravel()
    order_pred = np.argsort(-y_pred, kind="stable")
    order_true = np.argsort(-y_true, kind="stable")
    dcg = idcg = 0.0
    for i, j in enumerate(order_true[:k]):
        if y_true[j] > threshold:
            idcg += (2.0 ** y_true[j] - 1.0) / np.log(2.0 + i)
    for i, j in enumerate(order_pred[:k]):
        if y_true[j] > threshold:
            dcg += (2.0 ** y_true[j] - 1.0) / np.log(2.0 + i)
    return 0.0 if idcg == 0.0 else dcg / idcg


def mean_average_precision(y_true, y_pred, threshold=0.0):
    """Average precision of one query (Ranker.scala:152-172)."""
    y_true, y_pred = np.asarray(y_true, dtype=np.float64).ravel(), np.asarray(y_pred, dtype=np.float64).ravel()
    order = np.argsort(-y_pred, kind="stable")
    s, pos = 0.0, 0
    for i, j in enumerate(order):
        if y_true[j] > threshold:
            pos += 1
            s += pos / (i + 1.0)
    return 0.0 if pos == 0 else s / pos


class Ranker:
    """Mixin: `evaluate_ndcg` / `evaluate_map` over an iterable of (features, labels) query batches
    (a TextSet produced by `from_relation_lists`, or any list of pairs)."""

    def _score_queries(self, x):
        queries = x.to_query_batches() if hasattr(x, "to_query_batches") else x
        self.eval()
        dev = next(self.parameters()).device
        out = []
        with torch.no_grad():
            for feats, labels in queries:
                f = torch.as_tensor(np.asarray(feats)).to(dev)
                pred = self(f).float().reshape(-1).cpu().numpy()
                out.append((np.asarray(labels).reshape(-1), pred))
        return out

    def evaluate_ndcg(self, x, k, threshold=0.0):
        r = [ndcg(t, p, k, threshold) for t, p in self._score_queries(x)]
        return float(np.mean(r)) if r else 0.0

    def evaluate_map(self, x, threshold=0.0):
        r = [mean_average_precision(t, p, threshold) for t, p in self._score_queries(x)]
        return float(np.mean(r)) if r else 0.0
