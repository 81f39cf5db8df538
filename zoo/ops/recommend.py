"""Per-pair scoring and segmented top-N used by `Recommender.recommend_for_user/item` (kernel K15).

Reference: Recommender.scala:46-104 -- `predictUserItemPair` takes the arg-max class (1-based) and its
probability per pair; `recommendForUser` is `groupByKey(userId)` then
`sortBy(-prediction, -probability).take(maxItems)`.  Here: one scoring kernel builds a 64-bit
(class, probability) rank key, two radix sorts order the pairs by (key asc, rank desc), and one kernel
flags the first N rows of every key segment (`i < N or keys[i-N] != keys[i]`, no scan needed).
On CPU the same steps run in torch.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ..native import lib as _nl

_V, _I, _L = C.c_void_p, C.c_int, C.c_longlong
_nl.register_cuda_sigs({
    "zoo_pair_score": [_V, _L, _L, _I, _I, _V, _V, _V, _V],
    "zoo_segment_head": [_V, _L, _I, _V, _V],
})


def pair_score(probs: torch.Tensor):
    """probs [P, C] -> (class int32 1-based, probability fp32, rank_key int64)."""
    P, Cn = probs.shape
    if not probs.is_cuda:
        p, c = probs.float().max(-1)
        cls_ = (c + 1).int()
        key = (cls_.long() << 32) | p.clamp_min(0).view(torch.int32).long()
        return cls_, p, key
    if probs.dtype not in (torch.float32, torch.bfloat16):
        probs = probs.float()
    if probs.stride(-1) != 1:
        probs = probs.contiguous()
    cls_ = torch.empty(P, dtype=torch.int32, device=probs.device)
    p = torch.empty(P, dtype=torch.float32, device=probs.device)
    key = torch.empty(P, dtype=torch.int64, device=probs.device)
    _nl.check(_nl.cuda_lib().zoo_pair_score(probs.data_ptr(), probs.stride(0), P, Cn, 0 if probs.dtype == torch.float32 else 1,
                                            cls_.data_ptr(), p.data_ptr(), key.data_ptr(), _nl.stream_ptr()), "zoo_pair_score")
    return cls_, p, key


def segmented_topn_order(keys: torch.Tensor, rank_key: torch.Tensor, n: int):
    """Indices (into the inputs) of the top-`n` rows of every key, grouped by key asc, best first."""
    order = torch.argsort(rank_key, descending=True, stable=True)
    order = order[torch.argsort(keys[order], stable=True)]
    ks = keys[order].contiguous()
    if ks.is_cuda:
        keep = torch.empty(ks.numel(), dtype=torch.uint8, device=ks.device)
        _nl.check(_nl.cuda_lib().zoo_segment_head(ks.data_ptr(), ks.numel(), int(n), keep.data_ptr(), _nl.stream_ptr()),
                  "zoo_segment_head")
        keep = keep.bool()
    else:
        keep = torch.ones_like(ks, dtype=torch.bool)
        if ks.numel() > n:
            keep[n:] = ks[n:] != ks[:-n]
    return order[keep]


def segmented_topn(keys, cls_, prob, n, device=None):
    """numpy/torch inputs -> numpy indices of the top-`n` rows per key (sorted by -class, -probability)."""
    dev = torch.device(device) if device is not None else (torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu"))
    k = torch.as_tensor(np.asarray(keys), device=dev).long()
    c = torch.as_tensor(np.asarray(cls_), device=dev).long()
    p = torch.as_tensor(np.asarray(prob), device=dev).float()
    rank_key = (c << 32) | p.clamp_min(0).view(torch.int32).long()
    return segmented_topn_order(k, rank_key, n).cpu().numpy()
