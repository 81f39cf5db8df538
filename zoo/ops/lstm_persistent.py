"""Python face of the persistent recurrent kernels (K6, ``csrc/ops/rnn.cu``).

One kernel launch runs all T time steps of an LSTM / GRU layer (forward) and one more runs the
whole backward-through-time; the sequence-parallel parts -- the input projection
``x W_ih^T + b`` for all steps and the weight gradients, which are sums over B*T rows -- are plain
GEMMs issued around the kernels.  Weights use ``torch.nn.LSTM`` / ``torch.nn.GRU`` layout and gate
order (i,f,g,o / r,z,n), so the reference's PyTorch AutoML checkpoints
(pyzoo/zoo/automl/model/VanillaLSTM_pytorch.py:205-226) load unchanged.

Limits: unidirectional, batch-first, hidden size <= 128, fp32 I/O.  Everything else goes to
torch (cuDNN).  Backward consumes the saved gate buffer in place, so a graph can be
back-propagated once (no ``retain_graph``).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn.functional as F

from ..native import lib as _nl

_V, _I, _L = C.c_void_p, C.c_int, C.c_longlong
_nl.register_cuda_sigs({
    "zoo_rnn_supported": [_I],
    "zoo_lstm_forward": [_V, _V, _V, _V, _V, _V, _V, _V, _I, _I, _I, _I, _I, _V],
    "zoo_lstm_backward": [_V, _V, _V, _V, _V, _V, _V, _L, _L, _V, _V, _V, _V, _I, _I, _I, _I, _V],
    "zoo_gru_forward": [_V, _V, _V, _V, _V, _V, _V, _I, _I, _I, _I, _V],
    "zoo_gru_backward": [_V, _V, _V, _V, _V, _L, _L, _V, _V, _I, _I, _I, _I, _V],
})

MAX_HIDDEN = 128


def _p(t):
    return None if t is None else t.data_ptr()


def _sum_outer(dG, parts):
    """dG^T @ [parts..., 1] in ONE pass over dG: returns [G, sum(d_i) + 1] (the last column is the column sum of dG, i.e. the bias
    gradient).  The reduction runs over N = B*T rows with a tiny output, a shape library GEMMs handle badly without split-K
    (measured: 124 us per weight gradient at N 51,200, G 128, plus 125 us for the column sum); here N is cut into S slabs, a
    batched GEMM gives every slab its own partial product, and the slabs are summed."""
    N, G = dG.shape
    X = torch.cat(list(parts) + [torch.ones((N, 1), dtype=dG.dtype, device=dG.device)], dim=1)
    S = 1
    for cand in (256, 128, 64, 32, 16, 8, 4, 2):
        if N % cand == 0 and N // cand >= 64:
            S = cand
            break
    if S == 1:
        return dG.t() @ X
    return torch.bmm(dG.view(S, N // S, G).transpose(1, 2), X.view(S, N // S, X.shape[1])).sum(0)


def _dy_arg(dy):
    """(tensor kept alive, pointer, batch stride, time stride) for a [B,T,H] gradient."""
    if dy is None:
        return None, None, 0, 0
    if dy.dtype != torch.float32 or dy.stride(-1) != 1:
        dy = dy.float().contiguous()
    return dy, dy.data_ptr(), dy.stride(0), dy.stride(1)


_GRU_MASK = {}        # (H, device) -> [1]*2H + [0]*H: which entries of b_hh join the input bias


def _combined_bias(kind, b_ih, b_hh, H):
    """The bias the input projection adds, in ONE kernel: LSTM b_ih + b_hh; GRU b_ih with b_hh folded into the r and z gates only (the
    candidate gate's recurrent bias sits inside r * (.)).  Nothing derived from the parameters is cached: the fused optimizer writes
    parameters through raw pointers, which torch's version counters do not see."""
    if b_ih is None:
        return None
    if b_hh is None:
        return b_ih
    if kind == "lstm":
        return b_ih + b_hh
    key = (H, b_ih.device)
    mask = _GRU_MASK.get(key)
    if mask is None:
        mask = _GRU_MASK[key] = torch.cat([torch.ones(2 * H, device=b_ih.device), torch.zeros(H, device=b_ih.device)])
    return torch.addcmul(b_ih, b_hh, mask)


class _LSTMLayerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w_ih, w_hh, b_ih, b_hh, h0, c0, hard=False):
        B, T, Fin = x.shape
        H = w_hh.shape[1]
        x2 = x.reshape(B * T, Fin)
        need = any(ctx.needs_input_grad)
        bias = _combined_bias("lstm", b_ih, b_hh, H)
        gates = (torch.addmm(bias, x2, w_ih.t()) if bias is not None else x2 @ w_ih.t()).view(B, T, 4 * H)
        hout = torch.empty((B, T, H), dtype=torch.float32, device=x.device)
        cout = torch.empty_like(hout) if need else None
        hT = torch.empty((B, H), dtype=torch.float32, device=x.device)
        cT = torch.empty_like(hT)
        whh = w_hh.contiguous()
        _nl.check(_nl.cuda_lib().zoo_lstm_forward(gates.data_ptr(), whh.data_ptr(), _p(h0), _p(c0), hout.data_ptr(), _p(cout),
                                                  hT.data_ptr(), cT.data_ptr(), B, T, H, int(need), int(hard), _nl.stream_ptr()),
                  "zoo_lstm_forward")
        if need:
            ctx.save_for_backward(x, w_ih, whh, gates, hout, cout, h0, c0)
            ctx.has_bias = (b_ih is not None, b_hh is not None)
            ctx.hard = int(hard)
        return hout, hT, cT

    @staticmethod
    def backward(ctx, dy, dhT, dcT):
        x, w_ih, whh, gates, hout, cout, h0, c0 = ctx.saved_tensors
        B, T, Fin = x.shape
        H = whh.shape[1]
        dy_keep, dy_ptr, sb, st = _dy_arg(dy)
        dhT = None if dhT is None else dhT.float().contiguous()
        dcT = None if dcT is None else dcT.float().contiguous()
        dh0 = torch.empty((B, H), dtype=torch.float32, device=x.device) if ctx.needs_input_grad[5] else None
        dc0 = torch.empty((B, H), dtype=torch.float32, device=x.device) if ctx.needs_input_grad[6] else None
        _nl.check(_nl.cuda_lib().zoo_lstm_backward(gates.data_ptr(), whh.data_ptr(), hout.data_ptr(), cout.data_ptr(), _p(h0),
                                                   _p(c0), dy_ptr, sb, st, _p(dhT), _p(dcT), _p(dh0), _p(dc0), B, T, H,
                                                   ctx.hard, _nl.stream_ptr()), "zoo_lstm_backward")
        dG = gates.view(B * T, 4 * H)           # now holds the pre-activation gate gradients
        x2 = x.reshape(B * T, Fin)
        first = h0.unsqueeze(1) if h0 is not None else torch.zeros((B, 1, H), dtype=torch.float32, device=x.device)
        hprev = torch.cat([first, hout[:, :-1]], dim=1).reshape(B * T, H)
        dw_ih = dw_hh = db = None
        if any(ctx.needs_input_grad[1:5]):
            M = _sum_outer(dG, (x2, hprev))                   # [4H, Fin + H + 1]: both weight gradients and the bias gradient
            dw_ih, dw_hh, db = M[:, :Fin], M[:, Fin:Fin + H], M[:, Fin + H]
        dx = (dG @ w_ih).view(B, T, Fin) if ctx.needs_input_grad[0] else None
        return dx, dw_ih, dw_hh, (db if ctx.has_bias[0] else None), (db if ctx.has_bias[1] else None), dh0, dc0, None


class _GRULayerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w_ih, w_hh, b_ih, b_hh, h0, hard=False):
        B, T, Fin = x.shape
        H = w_hh.shape[1]
        x2 = x.reshape(B * T, Fin)
        need = any(ctx.needs_input_grad)
        bias = _combined_bias("gru", b_ih, b_hh, H)
        gx = (torch.addmm(bias, x2, w_ih.t()) if bias is not None else x2 @ w_ih.t()).view(B, T, 3 * H)
        hout = torch.empty((B, T, H), dtype=torch.float32, device=x.device)
        saved = torch.empty((B, T, 4 * H), dtype=torch.float32, device=x.device) if need else None
        hT = torch.empty((B, H), dtype=torch.float32, device=x.device)
        whh = w_hh.contiguous()
        bhn = None if b_hh is None else b_hh[2 * H:].contiguous()
        _nl.check(_nl.cuda_lib().zoo_gru_forward(gx.data_ptr(), _p(saved), whh.data_ptr(), _p(bhn), _p(h0), hout.data_ptr(),
                                                 hT.data_ptr(), B, T, H, int(hard), _nl.stream_ptr()), "zoo_gru_forward")
        if need:
            ctx.save_for_backward(x, w_ih, whh, saved, hout, h0)
            ctx.has_bias = (b_ih is not None, b_hh is not None)
            ctx.hard = int(hard)
        return hout, hT

    @staticmethod
    def backward(ctx, dy, dhT):
        x, w_ih, whh, saved, hout, h0 = ctx.saved_tensors
        B, T, Fin = x.shape
        H = whh.shape[1]
        dy_keep, dy_ptr, sb, st = _dy_arg(dy)
        dhT = None if dhT is None else dhT.float().contiguous()
        dh0 = torch.empty((B, H), dtype=torch.float32, device=x.device) if ctx.needs_input_grad[5] else None
        _nl.check(_nl.cuda_lib().zoo_gru_backward(saved.data_ptr(), whh.data_ptr(), hout.data_ptr(), _p(h0), dy_ptr, sb, st,
                                                  _p(dhT), _p(dh0), B, T, H, ctx.hard, _nl.stream_ptr()), "zoo_gru_backward")
        S = saved.view(B * T, 4 * H)            # dpr, dpz, dpn, dhn
        dGx = S[:, :3 * H]
        dRec = torch.cat([S[:, :2 * H], S[:, 3 * H:]], dim=1)
        x2 = x.reshape(B * T, Fin)
        first = h0.unsqueeze(1) if h0 is not None else torch.zeros((B, 1, H), dtype=torch.float32, device=x.device)
        hprev = torch.cat([first, hout[:, :-1]], dim=1).reshape(B * T, H)
        dw_ih = dw_hh = db_ih = db_hh = None
        if any(ctx.needs_input_grad[1:5]):
            M = _sum_outer(S, (x2, hprev))                    # [4H, Fin + H + 1] over the rows (dpr, dpz, dpn, dhn)
            rec = torch.cat([M[:2 * H], M[3 * H:]], dim=0)    # rows of the recurrent side: dpr, dpz, dhn
            dw_ih, dw_hh = M[:3 * H, :Fin], rec[:, Fin:Fin + H]
            db_ih = M[:3 * H, Fin + H] if ctx.has_bias[0] else None
            db_hh = rec[:, Fin + H] if ctx.has_bias[1] else None
        dx = (dGx @ w_ih).view(B, T, Fin) if ctx.needs_input_grad[0] else None
        return dx, dw_ih, dw_hh, db_ih, db_hh, dh0, None


def supported(rnn: torch.nn.Module, x: torch.Tensor) -> bool:
    return (x.is_cuda and x.dim() == 3 and isinstance(rnn, (torch.nn.LSTM, torch.nn.GRU)) and rnn.batch_first
            and not rnn.bidirectional and getattr(rnn, "proj_size", 0) == 0 and rnn.hidden_size <= MAX_HIDDEN)


def lstm_layer(x, w_ih, w_hh, b_ih=None, b_hh=None, h0=None, c0=None, hard_sigmoid=False):
    """One LSTM layer over a [B,T,F] fp32 sequence -> (outputs [B,T,H], h_T, c_T).
    `hard_sigmoid` selects the zoo keras default gate activation clip(0.2x+0.5, 0, 1)."""
    return _LSTMLayerFn.apply(x.float().contiguous(), w_ih, w_hh, b_ih, b_hh,
                              None if h0 is None else h0.float().contiguous(), None if c0 is None else c0.float().contiguous(),
                              bool(hard_sigmoid))


def gru_layer(x, w_ih, w_hh, b_ih=None, b_hh=None, h0=None, hard_sigmoid=False):
    return _GRULayerFn.apply(x.float().contiguous(), w_ih, w_hh, b_ih, b_hh, None if h0 is None else h0.float().contiguous(),
                             bool(hard_sigmoid))


def maybe_run(rnn, x, state=None, hard_sigmoid=False):
    """Run `rnn` (nn.LSTM / nn.GRU) on the persistent kernels; None when the module/shape is not covered."""
    if not supported(rnn, x):
        return None
    is_lstm = isinstance(rnn, torch.nn.LSTM)
    h0s = c0s = None
    if state is not None:
        h0s, c0s = state if is_lstm else (state, None)
    out, hs, cs = x, [], []
    for l in range(rnn.num_layers):
        g = lambda n: getattr(rnn, f"{n}_l{l}", None) if rnn.bias or n.startswith("weight") else None
        w_ih, w_hh, b_ih, b_hh = g("weight_ih"), g("weight_hh"), g("bias_ih"), g("bias_hh")
        h0 = None if h0s is None else h0s[l]
        if is_lstm:
            out, hT, cT = lstm_layer(out, w_ih, w_hh, b_ih, b_hh, h0, None if c0s is None else c0s[l], hard_sigmoid)
            cs.append(cT)
        else:
            out, hT = gru_layer(out, w_ih, w_hh, b_ih, b_hh, h0, hard_sigmoid)
        hs.append(hT)
        if rnn.dropout > 0 and l + 1 < rnn.num_layers and rnn.training:
            out = F.dropout(out, rnn.dropout, True)
    one = len(hs) == 1                                   # a single layer: the state is a view, no stacking kernel
    hN = hs[0].unsqueeze(0) if one else torch.stack(hs)
    return (out, (hN, cs[0].unsqueeze(0) if one else torch.stack(cs))) if is_lstm else (out, hN)
