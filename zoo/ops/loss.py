"""Fused classification losses (kernel K5, ``csrc/ops/loss.cu``).

`softmax_nll` = SoftMax + ZooClassNLLCriterion of the reference in one pass
(zoo/.../keras/objectives/ZooClassNLLCriterion.scala:48-121, SparseCategoricalCrossEntropy.scala:47-80):
probabilities are clipped to [eps, 1-eps] before the log, labels may be 0- or 1-based, rows whose
label equals `padding_value` are ignored, optional class weights, mean over the non-padding rows
when `size_average`.  `sigmoid_bce` is BinaryCrossEntropy.scala:36-42 on logits (or on
probabilities with `from_probs=True`).  The CPU path restates the same math with torch ops and is
the oracle of the GPU tests.
"""
from __future__ import annotations

import torch

from ..native import lib as _nl

EPS = 1e-8
# Set by the training engine around its own `loss.backward()`: the upstream gradient of the loss
# is exactly 1, so backward hands the kernel-produced dlogits straight on (no extra multiply).
UNIT_UPSTREAM = {"on": False}


def softmax_nll_reference(logits, labels, zero_based_label=True, padding_value=None, class_weights=None,
                          size_average=True, eps=EPS):
    """Plain-torch fp32 restatement (oracle + CPU path).  Returns (loss, probs)."""
    x = logits.float()
    p = torch.softmax(x, dim=-1)
    t = labels.long() - (0 if zero_based_label else 1)
    valid = (t >= 0) & (t < x.shape[-1])
    if padding_value is not None:
        valid = valid & (labels.long() != int(padding_value))
    tc = t.clamp(0, x.shape[-1] - 1)
    pt = p.gather(-1, tc.unsqueeze(-1)).squeeze(-1).clamp(eps, 1.0 - eps)
    w = class_weights.float()[tc] if class_weights is not None else torch.ones_like(pt)
    per = torch.where(valid, -w * torch.log(pt), torch.zeros_like(pt))
    if size_average:
        denom = valid.sum().clamp(min=1).to(per.dtype) if padding_value is not None else float(max(per.numel(), 1))
        return per.sum() / denom, p
    return per.sum(), p


class _SoftmaxNLLFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, zero_based, padding_value, class_w, size_average, eps):
        B, Cn = logits.shape
        lib = _nl.cuda_lib()
        dt = 0 if logits.dtype == torch.float32 else 1
        if logits.dtype not in (torch.float32, torch.bfloat16):
            logits, dt = logits.float(), 0
        if logits.stride(-1) != 1:
            logits = logits.contiguous()
        labels = labels.long().contiguous()
        out = torch.zeros(2, dtype=torch.float32, device=logits.device)
        ld = (Cn + 7) // 8 * 8  # padded rows: the next consumer is a TMA-fed GEMM (16-byte rows)
        dlog = torch.empty((B, ld), dtype=logits.dtype, device=logits.device)
        if ld != Cn:
            dlog[:, Cn:].zero_()
        dlog = dlog[:, :Cn]
        has_pad = padding_value is not None
        # mean over B is folded into the kernel; with padding the divisor is only known afterwards
        scale = (1.0 / max(B, 1)) if (size_average and not has_pad) else 1.0
        _nl.check(lib.zoo_softmax_nll(logits.data_ptr(), logits.stride(0), labels.data_ptr(), B, Cn, dt,
                                      0 if zero_based else 1, int(padding_value) if has_pad else 0, int(has_pad),
                                      _nl.ptr(class_w), float(eps), scale, scale, out.data_ptr(), None,
                                      dlog.data_ptr(), dlog.stride(0), None, 0, _nl.stream_ptr()), "zoo_softmax_nll")
        if size_average and has_pad:
            inv = 1.0 / out[1].clamp(min=1.0)
            loss = out[0] * inv
            dlog = dlog * inv.to(dlog.dtype)
        else:
            loss = out[0]
        ctx.save_for_backward(dlog)
        return loss

    @staticmethod
    def backward(ctx, g):
        (dlog,) = ctx.saved_tensors
        if UNIT_UPSTREAM["on"]:
            return dlog, None, None, None, None, None, None
        full = dlog.as_strided((dlog.shape[0], dlog.stride(0)), (dlog.stride(0), 1)) * g.to(dlog.dtype)
        return full[:, :dlog.shape[1]], None, None, None, None, None, None


def softmax_nll(logits, labels, zero_based_label=True, padding_value=None, class_weights=None, size_average=True,
                eps=EPS):
    """Scalar loss from raw logits [B, C] and integer labels [B]."""
    if logits.is_cuda:
        return _SoftmaxNLLFn.apply(logits, labels.reshape(-1), zero_based_label, padding_value, class_weights,
                                   size_average, eps)
    return softmax_nll_reference(logits, labels.reshape(-1), zero_based_label, padding_value, class_weights,
                                 size_average, eps)[0]


def sigmoid_bce_reference(x, y, from_probs=False, size_average=True, eps=EPS):
    p = x.float() if from_probs else torch.sigmoid(x.float())
    y = y.float()
    # p and 1-p are floored separately: 1 - 1e-8 rounds to 1 in fp32 (would give 0 * log(0) = nan)
    per = -(y * torch.log(p.clamp_min(eps)) + (1.0 - y) * torch.log((1.0 - p).clamp_min(eps)))
    return per.mean() if size_average else per.sum()


class _SigmoidBCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y, from_probs, size_average, eps):
        lib = _nl.cuda_lib()
        xs = x.contiguous()
        dt = 0 if xs.dtype == torch.float32 else 1
        if xs.dtype not in (torch.float32, torch.bfloat16):
            xs, dt = xs.float(), 0
        ys = y.float().contiguous()
        n = xs.numel()
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
        dx = torch.empty_like(xs)
        scale = 1.0 / max(n, 1) if size_average else 1.0
        _nl.check(lib.zoo_sigmoid_bce(xs.data_ptr(), ys.data_ptr(), n, dt, int(from_probs), float(eps), scale, scale,
                                      out.data_ptr(), dx.data_ptr(), _nl.stream_ptr()), "zoo_sigmoid_bce")
        ctx.save_for_backward(dx)
        ctx.shape = x.shape
        return out[0]

    @staticmethod
    def backward(ctx, g):
        (dx,) = ctx.saved_tensors
        return (dx * g.to(dx.dtype)).reshape(ctx.shape), None, None, None, None


def sigmoid_bce(x, y, from_probs=False, size_average=True, eps=EPS):
    if x.is_cuda:
        return _SigmoidBCEFn.apply(x, y, from_probs, size_average, eps)
    return sigmoid_bce_reference(x, y, from_probs, size_average, eps)
