"""Temporal convolution (TCN) ops: face of kernel K7 (``csrc/ops/tcn.cu``).

`causal_conv1d` = dilated causal conv + bias + ReLU + dropout (+ residual add + ReLU) in ONE kernel;
backward recovers the ReLU/dropout masks from the saved output (no mask tensor), runs the same tap
kernel mirrored for d(x), and a register-tiled reduction kernel for d(w).  `TemporalBlock` /
`TemporalConvNet` follow Bai et al.'s TCN (two weight-normalised dilated causal convs per block,
1x1 down-sample on the residual when channels change) -- the forecaster the north-star adds next to
the reference's LSTM / Seq2Seq / MTNet models.  CPU uses the same math in torch.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..native import lib as _nl

_V, _I, _L, _F = C.c_void_p, C.c_int, C.c_longlong, C.c_float
_nl.register_cuda_sigs({
    "zoo_tcn_conv_forward": [_V, _V, _V, _V, _V, _I, _I, _I, _I, _I, _I, _I, _F, C.c_ulonglong, _V],
    "zoo_tcn_conv_dgrad": [_V, _V, _V, _I, _I, _I, _I, _I, _I, _V],
    "zoo_tcn_prep_grad": [_V, _V, _V, _V, _V, _L, _I, _F, _V],
    "zoo_tcn_conv_wgrad": [_V, _V, _V, _I, _I, _I, _I, _I, _I, _V],
    # TF32 tensor-core forms (csrc/ops/tcn_mma.cu)
    "zoo_tcn_mma_supported": [_I, _I, _I, _I, _I],
    "zoo_tcn_mma_scratch_elems": [_I, _I, _I],
    "zoo_tcn_conv_forward_mma": [_V, _V, _V, _V, _V, _V, _I, _I, _I, _I, _I, _I, _I, _F, C.c_ulonglong, _V],
    "zoo_tcn_conv_dgrad_mma": [_V, _V, _V, _V, _I, _I, _I, _I, _I, _I, _V],
    "zoo_tcn_conv_wgrad_mma": [_V, _V, _V, _I, _I, _I, _I, _I, _I, _V],
})


def mma_path(Cin, Cout, T, K, dilation):
    """True when the TF32 tensor-core kernels take this layer (`ZOO_TCN_MMA=0` forces the fp32 CUDA-core kernels)."""
    import os

    if os.environ.get("ZOO_TCN_MMA", "1") == "0":
        return False
    return bool(_nl.cuda_lib().zoo_tcn_mma_supported(Cin, Cout, T, K, dilation))


def _mma_scratch(x, Cin, Cout, K):
    lib = _nl.cuda_lib()
    lib.zoo_tcn_mma_scratch_elems.restype = C.c_longlong
    return torch.empty(int(lib.zoo_tcn_mma_scratch_elems(Cin, Cout, K)), dtype=torch.float32, device=x.device)


def causal_conv1d_reference(x, w, bias=None, dilation=1, relu=True, res=None):
    """Plain torch restatement (no dropout): the CPU path and the oracle of the GPU tests."""
    K = w.shape[-1]
    y = F.conv1d(F.pad(x, ((K - 1) * dilation, 0)), w, bias, dilation=dilation)
    if relu:
        y = torch.relu(y)
    return torch.relu(y + res) if res is not None else y


class _CausalConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, res, dilation, relu, drop_p, seed):
        B, Cin, T = x.shape
        Cout, _, K = w.shape
        out = torch.empty((B, Cout, T), dtype=torch.float32, device=x.device)
        mma = mma_path(Cin, Cout, T, K, dilation)
        if mma:
            wt = _mma_scratch(x, Cin, Cout, K)
            _nl.check(_nl.cuda_lib().zoo_tcn_conv_forward_mma(x.data_ptr(), w.data_ptr(), wt.data_ptr(), _nl.ptr(bias), _nl.ptr(res),
                                                              out.data_ptr(), B, Cin, Cout, T, K, dilation, int(relu), float(drop_p),
                                                              int(seed), _nl.stream_ptr()), "zoo_tcn_conv_forward_mma")
        else:
            _nl.check(_nl.cuda_lib().zoo_tcn_conv_forward(x.data_ptr(), w.data_ptr(), _nl.ptr(bias), _nl.ptr(res), out.data_ptr(), B,
                                                          Cin, Cout, T, K, dilation, int(relu), float(drop_p), int(seed),
                                                          _nl.stream_ptr()), "zoo_tcn_conv_forward")
        ctx.save_for_backward(x, w, res, out)
        ctx.cfg = (dilation, relu, drop_p, bias is not None, mma)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, w, res, out = ctx.saved_tensors
        dilation, relu, drop_p, has_bias, mma = ctx.cfg
        B, Cin, T = x.shape
        Cout, _, K = w.shape
        lib, st = _nl.cuda_lib(), _nl.stream_ptr()
        dout = dout.float().contiguous()
        g = torch.empty_like(out)
        dres = torch.empty_like(out) if (res is not None and ctx.needs_input_grad[3]) else None
        _nl.check(lib.zoo_tcn_prep_grad(dout.data_ptr(), out.data_ptr(), _nl.ptr(res), g.data_ptr(), _nl.ptr(dres), out.numel(),
                                        int(relu), float(drop_p), st), "zoo_tcn_prep_grad")
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            if mma:
                wt = _mma_scratch(x, Cin, Cout, K)
                _nl.check(lib.zoo_tcn_conv_dgrad_mma(g.data_ptr(), w.data_ptr(), wt.data_ptr(), dx.data_ptr(), B, Cin, Cout, T, K, dilation, st),
                          "zoo_tcn_conv_dgrad_mma")
            else:
                _nl.check(lib.zoo_tcn_conv_dgrad(g.data_ptr(), w.data_ptr(), dx.data_ptr(), B, Cin, Cout, T, K, dilation, st),
                          "zoo_tcn_conv_dgrad")
        dw = None
        if ctx.needs_input_grad[1]:
            dw = torch.zeros_like(w)
            fn = lib.zoo_tcn_conv_wgrad_mma if mma else lib.zoo_tcn_conv_wgrad
            _nl.check(fn(g.data_ptr(), x.data_ptr(), dw.data_ptr(), B, Cin, Cout, T, K, dilation, st), "zoo_tcn_conv_wgrad")
        db = g.sum((0, 2)) if (has_bias and ctx.needs_input_grad[2]) else None
        return dx, dw, db, dres, None, None, None, None


def causal_conv1d(x, w, bias=None, dilation=1, relu=True, res=None, drop_p=0.0, training=False, seed=None):
    """x [B,Cin,T], w [Cout,Cin,K] -> [B,Cout,T]; see the module docstring for the fused epilogue."""
    p = float(drop_p) if training else 0.0
    if x.is_cuda:
        if seed is None:
            seed = int(torch.randint(0, 2**31 - 1, (1,)).item()) if p > 0 else 0
        return _CausalConvFn.apply(x.float().contiguous(), w.float().contiguous(), None if bias is None else bias.float(),
                                   None if res is None else res.float().contiguous(), int(dilation), bool(relu), p, seed)
    K = w.shape[-1]
    y = F.conv1d(F.pad(x, ((K - 1) * dilation, 0)), w, bias, dilation=dilation)
    if relu:
        y = torch.relu(y)
    if p > 0:
        y = F.dropout(y, p, True)
    return torch.relu(y + res) if res is not None else y


class CausalConv1d(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size, dilation=1, weight_norm=True, relu=True, dropout=0.0):
        super().__init__()
        self.conv = nn.Conv1d(in_channels, out_channels, kernel_size, dilation=dilation)
        nn.init.normal_(self.conv.weight, 0.0, 0.01)
        if weight_norm:
            self.conv = nn.utils.parametrizations.weight_norm(self.conv)
        self.dilation, self.relu, self.dropout = dilation, relu, dropout

    def forward(self, x, res=None):
        return causal_conv1d(x, self.conv.weight, self.conv.bias, self.dilation, self.relu, res, self.dropout, self.training)


class TemporalBlock(nn.Module):
    def __init__(self, n_in, n_out, kernel_size, dilation, dropout=0.2, weight_norm=True):
        super().__init__()
        self.conv1 = CausalConv1d(n_in, n_out, kernel_size, dilation, weight_norm, True, dropout)
        self.conv2 = CausalConv1d(n_out, n_out, kernel_size, dilation, weight_norm, True, dropout)
        self.downsample = nn.Conv1d(n_in, n_out, 1) if n_in != n_out else None
        if self.downsample is not None:
            nn.init.normal_(self.downsample.weight, 0.0, 0.01)

    def forward(self, x):
        if self.downsample is None:
            res = x
        else:  # 1x1 conv = a K=1 causal conv without activation
            res = causal_conv1d(x, self.downsample.weight, self.downsample.bias, 1, relu=False)
        return self.conv2(self.conv1(x), res=res)


class TemporalConvNet(nn.Module):
    """Stack of TemporalBlocks with dilation 2^i.  Input/outputs are [B, C, T]."""

    def __init__(self, num_inputs, num_channels, kernel_size=3, dropout=0.2, weight_norm=True):
        super().__init__()
        blocks, prev = [], num_inputs
        for i, ch in enumerate(num_channels):
            blocks.append(TemporalBlock(prev, ch, kernel_size, 2 ** i, dropout, weight_norm))
            prev = ch
        self.network = nn.Sequential(*blocks)

    def forward(self, x):
        return self.network(x)
