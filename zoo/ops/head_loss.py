"""Classification head in one kernel (K5b, ``csrc/ops/head_loss.cu``): the last Dense layer, softmax, NLL and the head's whole
backward (input gradient with the ReLU mask, weight and bias gradients) in one pass over the head's input.

Reference semantics: Dense(class_num) + SoftMax at the end of NeuralCF / Wide&Deep (zoo/.../models/recommendation/
NeuralCF.scala:89-96, WideAndDeep.scala:146-157) under ZooClassNLLCriterion / SparseCategoricalCrossEntropy
(ZooClassNLLCriterion.scala:48-121).  Used by the training engine only (`ENGINE_STEP`): the kernel accumulates the weight and bias
gradients into the flat gradient arena during FORWARD, which is right only when a backward with upstream gradient 1 follows and
the arena was zeroed by the previous optimizer step -- exactly the engine's step.  Everywhere else the layer-by-layer path runs.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from ..native import lib as _nl
from . import loss as _loss

_nl.register_cuda_sigs({
    "zoo_head_softmax_nll": [C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                             C.c_longlong, C.c_int, C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong,
                             C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_void_p],
})

# Set by the training engine around forward + backward of its own step (see module docstring).
ENGINE_STEP = {"on": False}


def head_reference(x, weight, bias, labels, zero_based=True, padding_value=None, size_average=True, relu_cols=0, eps=_loss.EPS):
    """Plain-torch fp32 restatement: (loss, dx, dw, db, logits).  The oracle of the GPU test."""
    x32 = x.float().detach().requires_grad_(True)
    w32 = weight.float().detach().requires_grad_(True)
    b32 = None if bias is None else bias.float().detach().requires_grad_(True)
    logits = x32 @ w32.t() + (0 if b32 is None else b32)
    loss, _ = _loss.softmax_nll_reference(logits, labels.reshape(-1), zero_based, padding_value, None, size_average, eps)
    grads = torch.autograd.grad(loss, [x32, w32] + ([b32] if b32 is not None else []))
    dx = grads[0]
    if relu_cols:
        dx = dx.clone()
        dx[:, :relu_cols] = torch.where(x32[:, :relu_cols] > 0, dx[:, :relu_cols], torch.zeros_like(dx[:, :relu_cols]))
    return loss.detach(), dx, grads[1], (grads[2] if b32 is not None else None), logits.detach()


def head_softmax_nll(x16, w16, bias, labels, dw32, db32, zero_based=True, padding_value=None, size_average=True, relu_cols=0,
                     eps=_loss.EPS, dx_out=None, want_logits=False, impl=None):
    """Raw launch.  x16 [B, K] bf16; w16 [C, >=K] bf16; dw32 [C, K] / db32 [C] fp32 are accumulated.  Returns (loss[1], dx, logits)."""
    B, K = x16.shape
    Cn = w16.shape[0]
    dev = x16.device
    out = torch.zeros(2, dtype=torch.float32, device=dev)          # [loss, valid-row count]
    dx = dx_out if dx_out is not None else torch.empty((B, K), dtype=torch.bfloat16, device=dev)
    logits = torch.empty((B, Cn), dtype=torch.float32, device=dev) if want_logits else None
    has_pad = padding_value is not None
    if impl is None:            # 0: warp-MMA kernel; 1: SIMT kernel (ZOO_HEAD_MMA=0)
        impl = 1 if os.environ.get("ZOO_HEAD_MMA", "1") == "0" else 0
    _nl.check(_nl.cuda_lib().zoo_head_softmax_nll(
        x16.data_ptr(), x16.stride(0), w16.data_ptr(), w16.stride(0), _nl.ptr(bias), labels.data_ptr(), B, K, Cn, int(bool(zero_based)),
        int(padding_value) if has_pad else 0, int(has_pad), int(bool(size_average)), float(eps), int(relu_cols), out.data_ptr(),
        dx.data_ptr(), dx.stride(0), dw32.data_ptr(), dw32.stride(0), _nl.ptr(db32), _nl.ptr(logits), Cn,
        out.data_ptr() + 4, int(impl), _nl.stream_ptr()), "zoo_head_softmax_nll")
    if has_pad and size_average:
        _nl.LAUNCH_COUNT["n"] += 1          # the valid-row count runs as its own small kernel first
    return out, dx, logits


class _HeadLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, w16, labels, zero_based, padding_value, size_average, relu_cols, dx_out):
        out, dx, _ = head_softmax_nll(x, w16, bias, labels, weight._zoo_grad, None if bias is None else bias._zoo_grad, zero_based,
                                      padding_value, size_average, relu_cols, dx_out=dx_out)
        ctx.save_for_backward(dx)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        (dx,) = ctx.saved_tensors
        if not _loss.UNIT_UPSTREAM["on"]:
            raise RuntimeError("the fused head is an engine-step path: its backward must be seeded with 1 (zoo.ops.loss.UNIT_UPSTREAM)")
        return dx, None, None, None, None, None, None, None, None, None


def applicable(x, layer, criterion, labels) -> bool:
    """True when (x -> layer -> criterion) can run as the one-pass head: engine step, CUDA bf16 input with 16-byte rows, a linear
    FusedDense with <= 8 outputs whose parameters live in the flat gradient arena, softmax-NLL without class weights."""
    from .dense import FusedDense

    if not (ENGINE_STEP["on"] and torch.is_grad_enabled() and isinstance(layer, FusedDense) and layer.activation is None):
        return False
    if not (torch.is_tensor(x) and x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.stride(1) == 1 and x.stride(0) % 8 == 0):
        return False
    K = x.shape[1]
    if K % 8 or K > 256 or K != layer.in_features or layer.out_features > 8 or x.data_ptr() % 16:
        return False
    if getattr(layer.weight, "_zoo_grad", None) is None or (layer.bias is not None and getattr(layer.bias, "_zoo_grad", None) is None):
        return False
    if getattr(criterion, "logits_kind", None) != "softmax" or getattr(criterion, "weights", None) is not None:
        return False
    if getattr(criterion, "log_prob_as_input", False):
        return False
    return torch.is_tensor(labels) and labels.is_cuda


def head_loss(x, layer, criterion, labels):
    """loss of criterion(softmax(layer(x)), labels) through the one-pass kernel.  Check `applicable(...)` first."""
    layer.refresh_shadow()
    y = labels.reshape(-1)
    if y.dtype != torch.int64:
        y = y.long()
    if not y.is_contiguous():
        y = y.contiguous()
    dx_out = getattr(x, "_zoo_dx_out", None)
    if dx_out is not None and (tuple(dx_out.shape) != tuple(x.shape) or dx_out.dtype != torch.bfloat16):
        dx_out = None
    relu_cols = (layer.relu_input_cols or layer.in_features) if layer.relu_input else 0
    return _HeadLossFn.apply(x, layer.weight, layer.bias, layer.weight_bf16, y, bool(criterion.zero_based_label), criterion.padding_value,
                             bool(criterion.size_average), int(relu_cols), dx_out)
