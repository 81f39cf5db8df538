"""Segmented top-N used by `Recommender.recommend_for_user/item` (kernel K15 on GPU).

Reference: Recommender.scala:46-58 -- `groupByKey(userId)` then
`sortBy(-prediction, -probability).take(maxItems)`.  Here: one stable lexicographic sort of
(key, -class, -probability) and a rank-within-segment filter, in torch (CUDA when available).
"""
from __future__ import annotations

import numpy as np
import torch


def segmented_topn(keys, cls_, prob, n, device=None):
    """Return indices (into the inputs) of the top-`n` rows of every key, grouped by key, best first."""
    dev = device or ("cuda" if torch.cuda.is_available() else "cpu")
    k = torch.as_tensor(np.asarray(keys), device=dev).long()
    c = torch.as_tensor(np.asarray(cls_), device=dev).long()
    p = torch.as_tensor(np.asarray(prob), device=dev).float()
    # stable sorts from the least significant key to the most significant one
    order = torch.argsort(p, descending=True, stable=True)
    order = order[torch.argsort(c[order], descending=True, stable=True)]
    order = order[torch.argsort(k[order], stable=True)]
    ks = k[order]
    start = torch.ones_like(ks, dtype=torch.bool)
    start[1:] = ks[1:] != ks[:-1]
    seg_start = torch.where(start, torch.arange(len(ks), device=dev), torch.zeros_like(ks)).cummax(0).values
    rank = torch.arange(len(ks), device=dev) - seg_start
    return order[rank < n].cpu().numpy()
