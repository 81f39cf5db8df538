"""Dense (Linear + bias + activation) on tcgen05 tensor cores.

`fused_dense(x, weight, bias, act)` runs the forward GEMM, the input-gradient GEMM (with the
ReLU mask of the producing layer fused into its epilogue) and the weight-gradient GEMM of
``csrc/gemm/dense_tcgen05.cu``.  On CPU (unit tests, gloo plumbing runs) the same math is done
with plain torch ops.

Replaces zoo keras ``Dense`` (reference zoo/.../pipeline/api/keras/layers/Dense.scala) and the
Linear+ReLU stacks of NeuralCF / WideAndDeep (NeuralCF.scala:73-96, WideAndDeep.scala:146-157).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..native import lib as _nl

import ctypes as _C

_nl.register_cuda_sigs({
    "zoo_mlp3_supported": [_C.c_int, _C.c_int, _C.c_int, _C.c_int],
    "zoo_mlp3_fwd_bf16": [_C.c_void_p, _C.c_longlong, _C.c_void_p, _C.c_longlong, _C.c_void_p, _C.c_void_p, _C.c_longlong, _C.c_void_p,
                          _C.c_void_p, _C.c_longlong, _C.c_void_p, _C.c_void_p, _C.c_longlong, _C.c_void_p, _C.c_longlong, _C.c_void_p,
                          _C.c_longlong, _C.c_int, _C.c_int, _C.c_int, _C.c_int, _C.c_int, _C.c_void_p, _C.c_void_p, _C.c_void_p, _C.c_void_p],
})

ACT_IDS = {None: 0, "linear": 0, "none": 0, "relu": 1, "gelu": 2, "sigmoid": 3, "tanh": 4}


def _act_torch(y, act):
    if act in (None, "linear", "none"):
        return y
    if act == "relu":
        return torch.relu(y)
    if act == "gelu":
        return F.gelu(y)
    if act == "sigmoid":
        return torch.sigmoid(y)
    if act == "tanh":
        return torch.tanh(y)
    raise ValueError(f"unsupported activation {act!r}")


def _use_kernel(x: torch.Tensor) -> bool:
    return x.is_cuda


def _aligned(t: torch.Tensor) -> bool:
    return t.stride(-1) == 1 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0


def dense_forward(x16: torch.Tensor, w16: torch.Tensor, bias: torch.Tensor | None, act, out: torch.Tensor | None = None,
                  relu_bits: torch.Tensor | None = None):
    """y = act(x16 @ w16.T + bias) with bf16 operands on the tcgen05 kernel.  `out` may be a
    column window of a wider buffer (fused concat).  `relu_bits` (int32 [M, ceil(N/32)]) optionally receives the
    sign bits of y for the consumer's dgrad (`dense_dgrad(mask_bits=...)`)."""
    M, K = x16.shape
    N = w16.shape[0]
    if out is None:
        ldn = (N + 7) // 8 * 8
        out = torch.empty((M, ldn), dtype=torch.bfloat16, device=x16.device)[:, :N]
    lib = _nl.cuda_lib()
    if not (_aligned(x16) and _aligned(w16) and _aligned(out)):
        raise _nl.NativeError("dense_forward: operands must have 16-byte aligned rows (pad K/N to a multiple of 8)")
    if relu_bits is not None:
        _nl.check(lib.zoo_dense_fwd_bf16_bits(x16.data_ptr(), x16.stride(0), w16.data_ptr(), w16.stride(0), _nl.ptr(bias),
                                              out.data_ptr(), out.stride(0), M, N, K, ACT_IDS[act], relu_bits.data_ptr(),
                                              relu_bits.stride(0), _nl.stream_ptr()), "zoo_dense_fwd_bf16_bits")
        return out
    _nl.check(lib.zoo_dense_fwd_bf16(x16.data_ptr(), x16.stride(0), w16.data_ptr(), w16.stride(0), _nl.ptr(bias),
                                     out.data_ptr(), out.stride(0), M, N, K, ACT_IDS[act], _nl.stream_ptr()),
              "zoo_dense_fwd_bf16")
    return out


def dense_dgrad(dy16, w16, mask16=None, out=None, mask_cols=0, mask_bits=None):
    """dx = (dy16 @ w16) * (mask16 > 0); with mask_cols > 0 only columns [0, mask_cols) are masked.  `mask_bits` is the
    bit-matrix form of the mask written by `dense_forward(relu_bits=...)`."""
    M, N = dy16.shape
    K = w16.shape[1]
    if out is None:
        ldk = (K + 7) // 8 * 8
        out = torch.empty((M, ldk), dtype=torch.bfloat16, device=dy16.device)[:, :K]
    lib = _nl.cuda_lib()
    if mask_bits is not None:
        _nl.check(lib.zoo_dense_dgrad_bf16_bits(dy16.data_ptr(), dy16.stride(0), w16.data_ptr(), w16.stride(0), mask_bits.data_ptr(),
                                                mask_bits.stride(0), int(mask_cols), out.data_ptr(), out.stride(0), M, N, K,
                                                _nl.stream_ptr()), "zoo_dense_dgrad_bf16_bits")
        return out
    _nl.check(lib.zoo_dense_dgrad_bf16(dy16.data_ptr(), dy16.stride(0), w16.data_ptr(), w16.stride(0),
                                       _nl.ptr(mask16), mask16.stride(0) if mask16 is not None else 0, int(mask_cols),
                                       out.data_ptr(), out.stride(0), M, N, K, _nl.stream_ptr()),
              "zoo_dense_dgrad_bf16")
    return out


def dense_wgrad(dy16, x16, dw32, db32=None):
    """dw32 += dy16.T @ x16  (fp32 accumulation straight into the gradient arena); with `db32` also db32 += dy16.sum(0),
    computed by the same kernel as one extra 16-column MMA against a tile of ones."""
    M, N = dy16.shape
    K = x16.shape[1]
    lib = _nl.cuda_lib()
    if db32 is not None and dw32.stride(0) % 4 == 0 and dw32.data_ptr() % 16 == 0:
        _nl.check(lib.zoo_dense_wgrad_bias_bf16(dy16.data_ptr(), dy16.stride(0), x16.data_ptr(), x16.stride(0), dw32.data_ptr(),
                                                dw32.stride(0), db32.data_ptr(), M, N, K, _nl.stream_ptr()), "zoo_dense_wgrad_bias_bf16")
        return dw32
    if db32 is not None:
        colsum(dy16, db32)
    if dw32.stride(0) % 4 != 0 or dw32.data_ptr() % 16 != 0:
        # the kernel reduces with 16-byte vector atomics: a weight whose row length is not a multiple of 4 floats
        # (e.g. in_features 75) is accumulated in a padded scratch and folded in
        ld = (K + 3) // 4 * 4
        tmp = torch.zeros((N, ld), dtype=torch.float32, device=dw32.device)
        _nl.check(lib.zoo_dense_wgrad_bf16(dy16.data_ptr(), dy16.stride(0), x16.data_ptr(), x16.stride(0),
                                           tmp.data_ptr(), tmp.stride(0), M, N, K, _nl.stream_ptr()), "zoo_dense_wgrad_bf16")
        dw32.add_(tmp[:, :K])
        return dw32
    _nl.check(lib.zoo_dense_wgrad_bf16(dy16.data_ptr(), dy16.stride(0), x16.data_ptr(), x16.stride(0),
                                       dw32.data_ptr(), dw32.stride(0), M, N, K, _nl.stream_ptr()),
              "zoo_dense_wgrad_bf16")
    return dw32


def colsum(dy16, db32):
    """db32 += dy16.sum(0)."""
    lib = _nl.cuda_lib()
    _nl.check(lib.zoo_colsum_bf16(dy16.data_ptr(), dy16.stride(0), dy16.shape[0], dy16.shape[1], db32.data_ptr(),
                                  _nl.stream_ptr()), "zoo_colsum_bf16")
    return db32


class _FusedDenseFn(torch.autograd.Function):
    """CUDA path.  Weight/bias gradients are accumulated straight into `.grad`-like fp32 buffers
    when the parameters carry ``_zoo_grad`` views (set by the flat-parameter arena)."""

    @staticmethod
    def forward(ctx, x, weight, bias, w16, act, relu_input, skip_relu_bwd=False, relu_input_cols=0, x_bits=None, emit_bits=False, dx_out=None,
                y_out=None):
        x16 = x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16)
        if not _aligned(x16):
            x16 = _pad_cols(x16)
        bits = None
        if emit_bits:
            bits = torch.empty((x16.shape[0], (w16.shape[0] + 31) // 32), dtype=torch.int32, device=x16.device)
        if y_out is not None:      # the caller's column window of a wider buffer (fused concat); returned as a fresh view
            y = dense_forward(x16, w16, bias, act, out=y_out, relu_bits=bits)[:, :]
        else:
            y = dense_forward(x16, w16, bias, act, relu_bits=bits)
        ctx.x_bits = x_bits
        ctx.dx_out = dx_out
        ctx.act = act
        ctx.relu_input = relu_input
        ctx.skip_relu_bwd = skip_relu_bwd
        ctx.relu_input_cols = relu_input_cols
        ctx.in_dtype = x.dtype
        ctx.save_for_backward(x16, w16, y if act == "relu" and not skip_relu_bwd else None, weight, bias)
        if emit_bits:
            ctx.mark_non_differentiable(bits)
            ctx.set_materialize_grads(False)      # no zero tensor for the bits output's "gradient"
            return y, bits
        return y

    @staticmethod
    def backward(ctx, dy, *_unused):
        x16, w16, y, weight, bias = ctx.saved_tensors
        if dy is None:
            return (None,) * 12
        dy16 = dy if dy.dtype == torch.bfloat16 else dy.to(torch.bfloat16)
        if ctx.act == "relu":
            if not ctx.skip_relu_bwd:  # otherwise the consumer's dgrad epilogue already applied the mask
                dy16 = torch.where(y > 0, dy16, torch.zeros_like(dy16))
        elif ctx.act not in (None, "linear", "none"):
            raise NotImplementedError("fused backward supports linear/relu epilogues; use torch for others")
        if not _aligned(dy16):
            dy16 = _pad_cols(dy16)
        K = weight.shape[1]
        dx = None
        if ctx.needs_input_grad[0]:
            wk = w16[:, :K] if w16.shape[1] != K else w16
            dxo = ctx.dx_out       # a buffer the producer of x wants its gradient in (e.g. peer-readable staging memory)
            if dxo is not None and (tuple(dxo.shape) != (dy16.shape[0], K) or dxo.dtype != torch.bfloat16 or ctx.in_dtype != torch.bfloat16):
                dxo = None
            if ctx.relu_input and ctx.x_bits is not None:
                dx = dense_dgrad(dy16, wk, out=dxo, mask_bits=ctx.x_bits, mask_cols=ctx.relu_input_cols)
            else:
                dx = dense_dgrad(dy16, wk, x16[:, :K] if ctx.relu_input else None, out=dxo, mask_cols=ctx.relu_input_cols)
            if dx.dtype != ctx.in_dtype:
                dx = dx.to(ctx.in_dtype)
        gw = getattr(weight, "_zoo_grad", None)
        gb = getattr(bias, "_zoo_grad", None) if bias is not None else None
        dw_ret, db_ret = None, None
        if bias is not None and gb is None and ctx.needs_input_grad[2]:
            gb = db_ret = torch.zeros_like(bias, dtype=torch.float32)
        if gw is not None:
            dense_wgrad(dy16, x16[:, :K], gw, gb)          # bias gradient rides along in the same kernel
        elif ctx.needs_input_grad[1]:
            dw_ret = torch.zeros_like(weight, dtype=torch.float32)
            dense_wgrad(dy16, x16[:, :K], dw_ret, gb)
        elif gb is not None:
            colsum(dy16, gb)
        return dx, dw_ret, db_ret, None, None, None, None, None, None, None, None, None


def _pad_cols(t: torch.Tensor) -> torch.Tensor:
    """Copy into a buffer whose row stride is a multiple of 8 elements (TMA needs 16-byte rows)."""
    M, K = t.shape
    ld = (K + 7) // 8 * 8
    buf = torch.zeros((M, ld), dtype=t.dtype, device=t.device)
    buf[:, :K] = t
    return buf[:, :K]


class FusedDense(nn.Module):
    """Linear + bias + activation.  fp32 master weights (optimizer state lives in the flat arena),
    bf16 shadow weights for the tensor-core GEMMs, refreshed by the fused optimizer kernel or
    lazily here when the master changes."""

    def __init__(self, in_features, out_features, activation=None, bias=True, init="glorot_uniform",
                 relu_input=False, relu_input_cols=0):
        super().__init__()
        self.in_features, self.out_features = int(in_features), int(out_features)
        self.activation = activation if activation not in ("linear", "none") else None
        self.relu_input = relu_input  # input is the output of a ReLU: fuse its mask into dgrad
        self.relu_input_cols = int(relu_input_cols)  # >0: only the first columns of the input are ReLU outputs
        self.skip_relu_bwd = False    # set when the ONLY consumer is a FusedDense with relu_input=True
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.zeros(out_features)) if bias else None
        self.register_buffer("weight_bf16", torch.zeros(out_features, (in_features + 7) // 8 * 8, dtype=torch.bfloat16),
                             persistent=False)
        self._shadow_version = -1
        self.reset_parameters(init)

    def reset_parameters(self, init="glorot_uniform"):
        if init in ("glorot_uniform", "xavier"):
            nn.init.xavier_uniform_(self.weight)
        elif init == "uniform":
            nn.init.uniform_(self.weight, -0.05, 0.05)
        elif init == "normal":
            nn.init.normal_(self.weight, 0.0, 0.05)
        else:
            nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        self._shadow_version = -1

    def refresh_shadow(self, force=False):
        v = self.weight._version
        if force or v != self._shadow_version or self.weight_bf16.device != self.weight.device:
            with torch.no_grad():
                self.weight_bf16[:, : self.in_features].copy_(self.weight)
            self._shadow_version = self.weight._version

    def mark_shadow_fresh(self):
        """Called by the fused optimizer after it wrote the bf16 shadow itself."""
        self._shadow_version = self.weight._version

    def forward(self, x, out=None):
        """`out` (CUDA path only): a bf16 [M, out_features] column window of a wider buffer that receives the result."""
        if _use_kernel(x):
            lead = x.shape[:-1]
            x2 = x.reshape(-1, x.shape[-1])
            self.refresh_shadow()
            # ReLU sign bits travel with the activation tensor: the producer writes them in its forward epilogue and
            # this layer's dgrad reads them instead of the full activation
            x_bits = getattr(x, "_zoo_relu_bits", None) if self.relu_input else None
            if x_bits is not None and (x_bits.shape[0] != x2.shape[0] or (self.relu_input_cols or self.in_features) % 32
                                       or x_bits.shape[1] * 32 < (self.relu_input_cols or self.in_features)):
                x_bits = None
            emit = self.activation == "relu" and self.skip_relu_bwd and torch.is_grad_enabled() and self.out_features % 32 == 0
            y = _FusedDenseFn.apply(x2, self.weight, self.bias, self.weight_bf16, self.activation, self.relu_input,
                                    self.skip_relu_bwd, self.relu_input_cols, x_bits, emit, getattr(x, "_zoo_dx_out", None), out)
            bits = None
            if emit:
                y, bits = y
            out = y.reshape(*lead, self.out_features)
            if bits is not None:
                out._zoo_relu_bits = bits
            return out
        y = F.linear(x.to(self.weight.dtype), self.weight, self.bias)
        return _act_torch(y, self.activation)

    def extra_repr(self):
        return f"{self.in_features}, {self.out_features}, activation={self.activation}"


def _tower_shapes_ok(layers, x):
    """The whole-tower kernel (csrc/gemm/mlp3_tcgen05.cu) covers three ReLU layers K0 -> N1 -> N2 -> 64 with widths in steps of 64."""
    if len(layers) != 3 or not all(isinstance(l, FusedDense) and l.activation == "relu" and l.bias is not None for l in layers):
        return False
    if not (torch.is_tensor(x) and x.is_cuda and x.dim() == 2 and x.dtype == torch.bfloat16 and _aligned(x)):
        return False
    if layers[0].in_features != x.shape[1] or layers[1].in_features != layers[0].out_features or layers[2].in_features != layers[1].out_features:
        return False
    try:
        return bool(_nl.cuda_lib().zoo_mlp3_supported(layers[0].in_features, layers[0].out_features, layers[1].out_features, layers[2].out_features))
    except AttributeError:
        return False


class _Tower3Fn(torch.autograd.Function):
    """Forward: one launch for the three layers (activations h1, h2, h3 and their ReLU sign bits are written for backward).
    Backward: the per-layer dgrad / wgrad kernels, exactly what three chained `_FusedDenseFn` would run."""

    @staticmethod
    def forward(ctx, x, w1, b1, w16_1, w2, b2, w16_2, w3, b3, w16_3, y_out, mask_last, relu_input, relu_input_cols, x_bits, dx_out):
        M, K0 = x.shape
        N1, N2, N3 = w16_1.shape[0], w16_2.shape[0], w16_3.shape[0]
        dev = x.device
        h1 = torch.empty((M, N1), dtype=torch.bfloat16, device=dev)
        h2 = torch.empty((M, N2), dtype=torch.bfloat16, device=dev)
        h3 = y_out if y_out is not None else torch.empty((M, N3), dtype=torch.bfloat16, device=dev)
        need = any(ctx.needs_input_grad)
        bits1 = torch.empty((M, N1 // 32), dtype=torch.int32, device=dev) if need else None
        bits2 = torch.empty((M, N2 // 32), dtype=torch.int32, device=dev) if need else None
        bits3 = torch.empty((M, N3 // 32), dtype=torch.int32, device=dev) if need else None
        _nl.check(_nl.cuda_lib().zoo_mlp3_fwd_bf16(
            x.data_ptr(), x.stride(0), w16_1.data_ptr(), w16_1.stride(0), _nl.ptr(b1), w16_2.data_ptr(), w16_2.stride(0), _nl.ptr(b2),
            w16_3.data_ptr(), w16_3.stride(0), _nl.ptr(b3), h1.data_ptr(), h1.stride(0), h2.data_ptr(), h2.stride(0), h3.data_ptr(), h3.stride(0),
            M, K0, N1, N2, N3, _nl.ptr(bits1), _nl.ptr(bits2), _nl.ptr(bits3), _nl.stream_ptr()), "zoo_mlp3_fwd_bf16")
        y = h3[:, :] if y_out is not None else h3
        if need:
            ctx.save_for_backward(x, h1, h2, y if mask_last else None, bits1, bits2, w16_1, w16_2, w16_3, w1, b1, w2, b2, w3, b3)
            ctx.mask_last, ctx.relu_input, ctx.relu_input_cols, ctx.x_bits, ctx.dx_out = mask_last, relu_input, relu_input_cols, x_bits, dx_out
            ctx.mark_non_differentiable(bits3)
            ctx.set_materialize_grads(False)
            return y, bits3
        return y, None

    @staticmethod
    def backward(ctx, dy, *_unused):
        x, h1, h2, y, bits1, bits2, w16_1, w16_2, w16_3, w1, b1, w2, b2, w3, b3 = ctx.saved_tensors
        none = (None,) * 16
        if dy is None:
            return none
        dy16 = dy if dy.dtype == torch.bfloat16 else dy.to(torch.bfloat16)
        if ctx.mask_last:                       # otherwise the consumer's dgrad epilogue (or the head kernel) already applied the mask
            dy16 = torch.where(y > 0, dy16, torch.zeros_like(dy16))
        if not _aligned(dy16):
            dy16 = _pad_cols(dy16)

        def wgrad(d, inp, w, b):
            gw, gb = getattr(w, "_zoo_grad", None), getattr(b, "_zoo_grad", None)
            rw = rb = None
            if gb is None:
                gb = rb = torch.zeros_like(b, dtype=torch.float32)
            if gw is None:
                gw = rw = torch.zeros_like(w, dtype=torch.float32)
            dense_wgrad(d, inp, gw, gb)
            return rw, rb

        dh2 = dense_dgrad(dy16, w16_3, mask_bits=bits2)                 # masks carry the ReLU of the layer below
        gw3, gb3 = wgrad(dy16, h2, w3, b3)
        dh1 = dense_dgrad(dh2, w16_2, mask_bits=bits1)
        gw2, gb2 = wgrad(dh2, h1, w2, b2)
        dx = None
        if ctx.needs_input_grad[0]:
            dxo = ctx.dx_out
            if dxo is not None and (tuple(dxo.shape) != tuple(x.shape) or dxo.dtype != torch.bfloat16):
                dxo = None
            if ctx.relu_input and ctx.x_bits is not None:
                dx = dense_dgrad(dh1, w16_1, out=dxo, mask_bits=ctx.x_bits, mask_cols=ctx.relu_input_cols)
            else:
                dx = dense_dgrad(dh1, w16_1, x if ctx.relu_input else None, out=dxo, mask_cols=ctx.relu_input_cols)
        gw1, gb1 = wgrad(dh1, x, w1, b1)
        return (dx, gw1, gb1, None, gw2, gb2, None, gw3, gb3, None) + (None,) * 6


class FusedMLP(nn.Sequential):
    """Dense(+ReLU) tower; every hidden layer feeds the next layer's dgrad its ReLU mask."""

    def __init__(self, in_features, hidden, activation="relu", last_activation=None, init="glorot_uniform"):
        layers, prev, prev_relu = [], in_features, False
        for i, h in enumerate(hidden):
            act = activation if i < len(hidden) - 1 or last_activation is None else last_activation
            layers.append(FusedDense(prev, h, activation=act, init=init, relu_input=prev_relu))
            prev, prev_relu = h, act == "relu"
        for a, b in zip(layers[:-1], layers[1:]):
            a.skip_relu_bwd = b.relu_input
        super().__init__(*layers)

    def forward(self, x, out=None):
        """`out` (CUDA path only): a bf16 [M, last width] column window of a wider buffer that receives the result.  Three-layer
        ReLU towers with supported widths run as ONE kernel (`ZOO_FUSED_TOWER=0` keeps the layer-by-layer path)."""
        import os

        layers = list(self)
        if os.environ.get("ZOO_FUSED_TOWER", "1") != "0" and _tower_shapes_ok(layers, x) and (out is None or _aligned(out)):
            l1, l2, l3 = layers
            for l in layers:
                l.refresh_shadow()
            x_bits = getattr(x, "_zoo_relu_bits", None) if l1.relu_input else None
            y, bits = _Tower3Fn.apply(x, l1.weight, l1.bias, l1.weight_bf16, l2.weight, l2.bias, l2.weight_bf16, l3.weight, l3.bias, l3.weight_bf16,
                                      out, not l3.skip_relu_bwd, l1.relu_input, l1.relu_input_cols, x_bits, getattr(x, "_zoo_dx_out", None))
            if bits is not None:
                y._zoo_relu_bits = bits
            return y
        h = x
        for l in layers[:-1]:
            h = l(h)
        return layers[-1](h, out=out) if (out is not None and isinstance(layers[-1], FusedDense)) else layers[-1](h)
