"""FP8 (E4M3) inference for Linear / Dense layers: the low-precision inference engine of the serving path.

The reference gets its fast inference from 8-bit engines -- OpenVINO int8 IR and BigDL's int8 quantisation
(zoo/.../pipeline/inference/OpenVinoInferenceSupportive.scala:35-57, `InferenceModel.doLoadOpenVINOInt8`,
docs/docs/wp-bigdl.md:192-196: "< 0.1 % accuracy drop, 4x smaller, up to 2x faster").  On B200 the 8-bit tensor-core format is
FP8: `quantize_linear_layers(model, calib_inputs)` replaces `nn.Linear` / `FusedDense` layers by `FP8Linear`, whose GEMM is the
FP8 mode of the hand-written tcgen05 kernel (csrc/gemm/dense_tcgen05.cu: E4M3 operands, fp32 accumulation in TMEM, per-tensor
scales folded into the epilogue, bias + activation fused).  Weights are quantised once (per-tensor absmax scale); activation
scales come from a calibration batch, like the reference's calibration step for int8 IR.  Consecutive FP8 layers pass E4M3
activations to each other straight from the producing kernel's epilogue (no separate quantise pass, half the bytes of bf16).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from ..native import lib as _nl
from .dense import ACT_IDS

E4M3_MAX = 448.0
_nl.register_cuda_sigs({
    "zoo_dense_fwd_fp8": [C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_int,
                          C.c_int, C.c_float, C.c_int, C.c_float, C.c_void_p],
    "zoo_quantize_e4m3": [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_longlong, C.c_longlong, C.c_int, C.c_float, C.c_void_p],
    "zoo_absmax": [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p],
})


def _pad16(n):
    return (n + 15) // 16 * 16


def quantize_e4m3(x: torch.Tensor, scale: float, out: torch.Tensor | None = None):
    """x [M, K] (fp32 / bf16, CUDA) -> uint8 [M, pad16(K)] holding E4M3(x / scale); the padding columns are zero."""
    M, K = x.shape
    if x.dtype not in (torch.float32, torch.bfloat16):
        x = x.float()
    if x.stride(-1) != 1:
        x = x.contiguous()
    if out is None:
        out = torch.empty((M, _pad16(K)), dtype=torch.uint8, device=x.device)
    if x.is_cuda:
        _nl.check(_nl.cuda_lib().zoo_quantize_e4m3(x.data_ptr(), x.stride(0), 0 if x.dtype == torch.float32 else 1, out.data_ptr(), out.stride(0),
                                                   M, K, 1.0 / scale, _nl.stream_ptr()), "zoo_quantize_e4m3")
    else:   # CPU oracle: torch's own E4M3 conversion
        out.zero_()
        out[:, :K] = (x.float() / scale).clamp(-E4M3_MAX, E4M3_MAX).to(torch.float8_e4m3fn).view(torch.uint8)
    return out


def dequantize_e4m3(q: torch.Tensor, scale: float, cols=None):
    v = q.view(torch.float8_e4m3fn).float() * scale
    return v if cols is None else v[:, :cols]


def absmax(x: torch.Tensor) -> float:
    if x.is_cuda and x.dtype in (torch.float32, torch.bfloat16) and x.is_contiguous():
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
        _nl.check(_nl.cuda_lib().zoo_absmax(x.data_ptr(), x.numel(), 0 if x.dtype == torch.float32 else 1, out.data_ptr(), _nl.stream_ptr()), "zoo_absmax")
        return float(out)
    return float(x.detach().abs().max())


class FP8Tensor:
    """An activation that already is E4M3: bytes [M, pad16(K)] + the scale that maps them back to real values."""
    __slots__ = ("q", "scale", "cols")

    def __init__(self, q, scale, cols):
        self.q, self.scale, self.cols = q, float(scale), int(cols)

    def float(self):
        return dequantize_e4m3(self.q, self.scale, self.cols)


class FP8Linear(nn.Module):
    """y = act(x W^T + b) with E4M3 operands.  `in_scale`: quantisation scale of the input (from calibration);
    `next_scale`: when set, the output is emitted as an FP8Tensor quantised with that scale (the next FP8Linear's `in_scale`)."""

    def __init__(self, weight, bias, activation=None, in_scale=1.0, next_scale=None):
        super().__init__()
        w = weight.detach().float()
        self.out_features, self.in_features = w.shape
        self.activation = activation
        self.w_scale = max(float(w.abs().max()), 1e-12) / E4M3_MAX
        wq = torch.zeros((self.out_features, _pad16(self.in_features)), dtype=torch.uint8, device=w.device)
        wq[:, :self.in_features] = (w / self.w_scale).clamp(-E4M3_MAX, E4M3_MAX).to(torch.float8_e4m3fn).view(torch.uint8)
        self.register_buffer("wq", wq)
        self.register_buffer("bias", None if bias is None else bias.detach().float().clone())
        self.in_scale, self.next_scale = float(in_scale), next_scale

    def forward(self, x):
        if isinstance(x, FP8Tensor):
            xq, sx = x.q, x.scale
            lead = None
        else:
            lead = x.shape[:-1]
            x2 = x.reshape(-1, x.shape[-1])
            sx = self.in_scale
            xq = quantize_e4m3(x2, sx)
        M = xq.shape[0]
        if not xq.is_cuda:         # CPU oracle of the same arithmetic (dequantised operands, fp32 accumulation)
            y = dequantize_e4m3(xq, sx, self.in_features) @ dequantize_e4m3(self.wq, self.w_scale, self.in_features).t()
            if self.bias is not None:
                y = y + self.bias
            from .dense import _act_torch

            y = _act_torch(y, self.activation)
            if self.next_scale is not None:
                return FP8Tensor(quantize_e4m3(y, self.next_scale), self.next_scale, self.out_features)
            return y.reshape(*lead, self.out_features) if lead is not None else y
        lib = _nl.cuda_lib()
        if self.next_scale is not None:
            out = torch.empty((M, _pad16(self.out_features)), dtype=torch.uint8, device=xq.device)
            if out.shape[1] != self.out_features:
                out[:, self.out_features:].zero_()
            _nl.check(lib.zoo_dense_fwd_fp8(xq.data_ptr(), xq.stride(0), self.wq.data_ptr(), self.wq.stride(0), _nl.ptr(self.bias), out.data_ptr(),
                                            out.stride(0), M, self.out_features, self.wq.shape[1], ACT_IDS[self.activation], sx * self.w_scale, 1,
                                            1.0 / self.next_scale, _nl.stream_ptr()), "zoo_dense_fwd_fp8")
            return FP8Tensor(out, self.next_scale, self.out_features)
        ldn = (self.out_features + 7) // 8 * 8
        out = torch.empty((M, ldn), dtype=torch.bfloat16, device=xq.device)
        _nl.check(lib.zoo_dense_fwd_fp8(xq.data_ptr(), xq.stride(0), self.wq.data_ptr(), self.wq.stride(0), _nl.ptr(self.bias), out.data_ptr(),
                                        out.stride(0), M, self.out_features, self.wq.shape[1], ACT_IDS[self.activation], sx * self.w_scale, 0, 1.0,
                                        _nl.stream_ptr()), "zoo_dense_fwd_fp8")
        y = out[:, :self.out_features]
        return y.reshape(*lead, self.out_features) if lead is not None else y

    def extra_repr(self):
        return f"{self.in_features}, {self.out_features}, activation={self.activation}, fp8(e4m3), chained={self.next_scale is not None}"


def _linear_like(m):
    from .dense import FusedDense

    if isinstance(m, FusedDense):
        return m.weight, m.bias, m.activation
    if isinstance(m, nn.Linear):
        return m.weight, m.bias, None
    return None


@torch.no_grad()
def quantize_linear_layers(model: nn.Module, calib_inputs, skip=()):
    """Replace every `nn.Linear` / `FusedDense` of `model` (except names in `skip`) by an `FP8Linear`, in place.

    calib_inputs: one input (tensor or tuple of tensors) or a list of them; the model is run on it once to record the absmax of every
    layer's input (the activation scale).  Layers that directly follow each other inside an `nn.Sequential` (FusedMLP included) are
    chained: the first emits E4M3 for the second.  Returns the model (eval mode)."""
    model.eval()
    targets = {n: m for n, m in model.named_modules() if _linear_like(m) is not None and n not in skip}
    if not targets:
        return model
    amax = {n: 0.0 for n in targets}
    hooks = []
    for n, m in targets.items():
        hooks.append(m.register_forward_pre_hook(lambda mod, inp, n=n: amax.__setitem__(n, max(amax[n], absmax(inp[0].detach().contiguous())))))
    batches = calib_inputs if isinstance(calib_inputs, list) else [calib_inputs]
    for b in batches:
        model(*b) if isinstance(b, tuple) else model(b)
    for h in hooks:
        h.remove()
    new = {}
    for n, m in targets.items():
        w, b, act = _linear_like(m)
        new[n] = FP8Linear(w, b, act, in_scale=max(amax[n], 1e-12) / E4M3_MAX)
    # chain neighbours inside Sequential containers
    for cname, cont in model.named_modules():
        if isinstance(cont, nn.Sequential):
            kids = list(cont.named_children())
            for (na, _), (nb, _) in zip(kids[:-1], kids[1:]):
                fa, fb = (f"{cname}.{na}" if cname else na), (f"{cname}.{nb}" if cname else nb)
                if fa in new and fb in new:
                    new[fa].next_scale = new[fb].in_scale
    for n, q in new.items():
        parent = model
        parts = n.split(".")
        for p in parts[:-1]:
            parent = getattr(parent, p)
        setattr(parent, parts[-1], q)
    # a chain's last link may feed a non-FP8 consumer: containers unwrap FP8Tensor at their end
    for cont in model.modules():
        if isinstance(cont, nn.Sequential) and not getattr(cont, "_zoo_fp8_unwrap", False):
            cont.register_forward_hook(lambda mod, inp, out: out.float() if isinstance(out, FP8Tensor) else out)
            cont._zoo_fp8_unwrap = True
    return model
