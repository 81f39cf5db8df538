"""Scaled-dot-product attention used by TransformerLayer / BERT.

Reference: TransformerLayer.scala:163-181 materialises `[batch, heads, L, L]` scores (`mm(q,k)` -> mask ->
softmax -> dropout -> `mm(w,v)`).  Here one fused call is used: blockwise (flash-style) attention through
torch's SDPA on the GPU, which never writes the L x L matrix; the explicit reference math is kept as
`attention_reference` (CPU path and oracle).  A hand-written tcgen05 kernel for this op is the K8 stretch
item of SURVEY section 8; until it lands this op counts as a library call.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def attention_reference(q, k, v, additive_mask=None, causal=False, scale=True, dropout_p=0.0, training=False):
    """q,k,v: [B, heads, L, d].  `additive_mask` broadcasts to [B, heads, L, L] (0 keep / -1e4 drop)."""
    w = torch.matmul(q, k.transpose(-1, -2))
    if scale:
        w = w / math.sqrt(v.shape[-1])
    if causal:
        L = w.shape[-1]
        tri = torch.tril(torch.ones(L, L, device=w.device, dtype=w.dtype))
        w = w * tri + (1.0 - tri) * (-1e9)
    if additive_mask is not None:
        w = w + additive_mask
    w = torch.softmax(w, -1)
    if dropout_p > 0 and training:
        w = F.dropout(w, dropout_p, True)
    return torch.matmul(w, v)


def attention(q, k, v, additive_mask=None, causal=False, scale=True, dropout_p=0.0, training=False):
    if not q.is_cuda:
        return attention_reference(q, k, v, additive_mask, causal, scale, dropout_p, training)
    mask = None
    if additive_mask is not None:
        mask = additive_mask.to(q.dtype)
        if causal:
            L = q.shape[-2]
            tri = torch.ones(L, L, device=q.device, dtype=torch.bool).tril()
            mask = mask.masked_fill(~tri, -1e9) if mask.shape[-2] == L else (mask + torch.zeros(L, L, device=q.device, dtype=q.dtype)).masked_fill(~tri, -1e9)
            causal = False
    return F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=dropout_p if training else 0.0, is_causal=causal,
                                          scale=None if scale else 1.0)
