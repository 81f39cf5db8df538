"""Scaled-dot-product attention used by TransformerLayer / BERT.

Reference: TransformerLayer.scala:163-181 materialises `[batch, heads, L, L]` scores (`mm(q,k)` -> mask ->
softmax -> dropout -> `mm(w,v)`).  Here one fused call is used: blockwise (flash-style) attention through
torch's SDPA on the GPU, which never writes the L x L matrix; the explicit reference math is kept as
`attention_reference` (CPU path and oracle).  The hand-written tcgen05 kernels for this op (K8: csrc/ops/attention.cu forward, csrc/ops/attention_bwd.cu
backward with recomputed probabilities; head dimension 64, key-padding mask and causal) are the default on the GPU; the cases they do
not cover (other head dimensions, arbitrary [L, L] masks, attention dropout in training) take torch SDPA.
"""
from __future__ import annotations

import math
import os

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


def attention_blockwise_reference(q, k, v, key_mask=None, causal=False, scale=True, block=128):
    """The K8 kernel's algorithm restated in torch (CPU oracle for debugging it): 128-query tiles walk the keys in blocks of
    `block`; scores in the log2 domain, running maximum m, running sum l, output rescaled by alpha = 2^(m_old - m_new); the
    probabilities are rounded to bf16 before the second product, as the kernel feeds them to the tensor core; keys beyond L or
    above the diagonal get probability 0; a row with no live key yields zeros."""
    B, H, L, D = q.shape
    log2e = 1.4426950408889634
    sc = ((1.0 / math.sqrt(D)) if scale else 1.0) * log2e
    out = torch.zeros(B, H, L, D, dtype=torch.float32, device=q.device)
    qf, kf, vf = q.float(), k.float(), v.float()
    for q0 in range(0, L, 128):
        q1 = min(L, q0 + 128)
        rows = q1 - q0
        o = torch.zeros(B, H, rows, D, device=q.device)
        m = torch.full((B, H, rows), float("-inf"), device=q.device)
        l = torch.zeros(B, H, rows, device=q.device)
        qi = torch.arange(q0, q1, device=q.device)
        n_blocks = (L + block - 1) // block
        if causal:
            n_blocks = min(n_blocks, (q0 + 128 + block - 1) // block)
        for j in range(n_blocks):
            k0, k1 = j * block, min(L, (j + 1) * block)
            s = (qf[:, :, q0:q1] @ kf[:, :, k0:k1].transpose(-1, -2)) * sc
            if key_mask is not None:
                s = s + key_mask[:, None, None, k0:k1].float() * log2e
            if causal:
                key = torch.arange(k0, k1, device=q.device)
                s = s.masked_fill(key[None, :] > qi[:, None], float("-inf"))
            m_new = torch.maximum(m, s.max(-1).values)
            alpha = torch.where(torch.isinf(m) & (m < 0), torch.zeros_like(m), torch.exp2(m - m_new))
            m_use = torch.where(torch.isinf(m_new), torch.zeros_like(m_new), m_new)
            p = torch.exp2(s - m_use[..., None])
            l = l * alpha + p.sum(-1)
            o = o * alpha[..., None] + p.to(torch.bfloat16).float() @ vf[:, :, k0:k1]
            m = m_new
        out[:, :, q0:q1] = o * torch.where(l > 0, 1.0 / l, torch.zeros_like(l))[..., None]
    return out


class _AttentionFn(torch.autograd.Function):
    """K8: forward keeps one log-sum-exp per row; backward recomputes the probabilities block by block (csrc/ops/attention_bwd.cu)."""

    @staticmethod
    def forward(ctx, q, k, v, key_mask, causal, scale):
        from ..native import lib as _nl

        B, H, L, D = q.shape
        q16, k16, v16 = (t.to(torch.bfloat16).contiguous() for t in (q, k, v))
        out = torch.empty_like(q16)
        km = None if key_mask is None else key_mask.to(torch.float32).contiguous()
        sc = (1.0 / math.sqrt(D)) if scale else 1.0
        need_bwd = any(ctx.needs_input_grad[:3])
        lib = _nl.cuda_lib()
        if need_bwd:
            lse = torch.empty((B, H, L), dtype=torch.float32, device=q.device)
            _nl.check(lib.zoo_attention_fwd_lse_bf16(q16.data_ptr(), k16.data_ptr(), v16.data_ptr(), out.data_ptr(), lse.data_ptr(), _nl.ptr(km),
                                                     B, H, L, D, sc, int(causal), _nl.stream_ptr()), "zoo_attention_fwd_lse_bf16")
            ctx.save_for_backward(q16, k16, v16, out, lse, km)
            ctx.sc, ctx.causal, ctx.dtypes = sc, bool(causal), (q.dtype, k.dtype, v.dtype)
        else:
            _nl.check(lib.zoo_attention_fwd_bf16(q16.data_ptr(), k16.data_ptr(), v16.data_ptr(), out.data_ptr(), _nl.ptr(km), B, H, L, D, sc,
                                                 int(causal), _nl.stream_ptr()), "zoo_attention_fwd_bf16")
        return out.to(q.dtype)

    @staticmethod
    def backward(ctx, d_out):
        from ..native import lib as _nl

        q16, k16, v16, out, lse, km = ctx.saved_tensors
        B, H, L, D = q16.shape
        do16 = d_out.to(torch.bfloat16).contiguous()
        dq = torch.zeros((B, H, L, D), dtype=torch.float32, device=q16.device)       # key blocks accumulate into it
        dk, dv = torch.empty_like(k16), torch.empty_like(v16)
        delta = torch.empty((B, H, L), dtype=torch.float32, device=q16.device)
        _nl.check(_nl.cuda_lib().zoo_attention_bwd_bf16(q16.data_ptr(), k16.data_ptr(), v16.data_ptr(), out.data_ptr(), do16.data_ptr(),
                                                        lse.data_ptr(), delta.data_ptr(), _nl.ptr(km), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(),
                                                        B, H, L, D, ctx.sc, int(ctx.causal), _nl.stream_ptr()), "zoo_attention_bwd_bf16")
        _nl.LAUNCH_COUNT["n"] += 1      # delta pre-pass + the main kernel
        tq, tk, tv = ctx.dtypes
        return dq.to(tq), dk.to(tk), dv.to(tv), None, None, None


def attention_kernel(q, k, v, key_mask=None, causal=False, scale=True):
    """K8.  q, k, v: [B, H, L, 64] (any float dtype; computed in bf16); key_mask: additive [B, L] or None.  Differentiable."""
    return _AttentionFn.apply(q, k, v, key_mask, causal, scale)


def _kernel_eligible(q, additive_mask, dropout_p, training):
    """The kernel covers head dimension 64, key-padding masks and the causal mask, without attention dropout in training (the
    remaining cases take torch SDPA).  ZOO_ATTN_KERNEL=0 switches it off."""
    if os.environ.get("ZOO_ATTN_KERNEL", "1") == "0" or q.shape[-1] != 64 or (dropout_p > 0 and training):
        return False
    if additive_mask is None:
        return True
    m = additive_mask                     # only key-padding masks [B, 1, 1, L] (or [B, L]) map onto the kernel
    return (m.dim() == 4 and m.shape[1] == 1 and m.shape[2] == 1 and m.shape[0] == q.shape[0]) or (m.dim() == 2 and m.shape[0] == q.shape[0])


def attention(q, k, v, additive_mask=None, causal=False, scale=True, dropout_p=0.0, training=False):
    if not q.is_cuda:
        return attention_reference(q, k, v, additive_mask, causal, scale, dropout_p, training)
    if _kernel_eligible(q, additive_mask, dropout_p, training):
        km = None if additive_mask is None else additive_mask.reshape(additive_mask.shape[0], -1)
        return attention_kernel(q, k, v, km, causal, scale)
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
