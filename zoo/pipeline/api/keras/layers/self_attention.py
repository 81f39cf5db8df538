"""TransformerLayer and BERT (reference: layers/self_attention.py:46-400; Scala TransformerLayer.scala:56-206,
BERT.scala).  Same block structure as the reference -- fused QKV projection, multi-head attention (causal
mask when not bidirectional, additive attention mask for BERT), post-LayerNorm, GELU MLP (tanh approximation
for TransformerLayer, erf for BERT), tanh pooler over the first token -- written as modules: the
projections are tcgen05 Dense layers (zoo.ops.dense), attention is one fused call (zoo.ops.attention)
instead of a materialised L x L softmax."""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .....ops.attention import attention
from .....ops.dense import FusedDense
from .base import ZooKerasLayer


def _gelu_tanh(x):
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x * x * x)))


def _gelu_erf(x):
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


class TransformerBlock(nn.Module):
    def __init__(self, hidden, n_head, intermediate, hidden_drop, attn_drop, init_range, causal, gelu, ln_eps=1e-5):
        super().__init__()
        if hidden % n_head:
            raise ValueError(f"hidden size {hidden} is not a multiple of n_head {n_head}")
        self.hidden, self.n_head, self.causal = hidden, n_head, causal
        self.qkv = FusedDense(hidden, 3 * hidden, init="normal")
        self.proj = FusedDense(hidden, hidden, init="normal")
        self.fc1 = FusedDense(hidden, intermediate, init="normal")
        self.fc2 = FusedDense(intermediate, hidden, init="normal")
        for m in (self.qkv, self.proj, self.fc1, self.fc2):
            nn.init.normal_(m.weight, 0.0, init_range)
        self.ln1, self.ln2 = nn.LayerNorm(hidden, eps=ln_eps), nn.LayerNorm(hidden, eps=ln_eps)
        self.hidden_drop, self.attn_drop, self.gelu = hidden_drop, attn_drop, gelu

    def forward(self, x, mask=None):
        B, L, H = x.shape
        d = H // self.n_head
        qkv = self.qkv(x).reshape(B, L, 3, self.n_head, d).permute(2, 0, 3, 1, 4)
        a = attention(qkv[0], qkv[1], qkv[2], mask, causal=self.causal, scale=True, dropout_p=self.attn_drop, training=self.training)
        a = self.proj(a.transpose(1, 2).reshape(B, L, H))
        a = F.dropout(a.float(), self.hidden_drop, self.training)
        n = self.ln1(x.float() + a)
        m = self.fc2(self.gelu(self.fc1(n).float()))
        m = F.dropout(m.float(), self.hidden_drop, self.training)
        return self.ln2(n + m)


class TransformerLayer(ZooKerasLayer):
    """Inputs [token ids (B,L), position ids (B,L)] -> [last block output (B,L,H), pooled first token (B,H)]
    (or every block's output + pooled with `output_all_block`)."""
    multi_input = True
    input_dtype = torch.int64

    def __init__(self, n_block, hidden_drop, attn_drop, n_head, initializer_range, bidirectional, output_all_block,
                 embedding_layer, input_shape, intermediate_size=0, hidden_size=None, **kw):
        super().__init__(input_shape, **kw)
        self.n_block, self.hidden_drop, self.attn_drop, self.n_head = int(n_block), hidden_drop, attn_drop, int(n_head)
        self.initializer_range, self.bidirectional, self.output_all_block = initializer_range, bidirectional, output_all_block
        self.embedding_layer, self.intermediate_size = embedding_layer, intermediate_size
        self.seq_len = input_shape[0][0]
        self._hidden = hidden_size
        self.ensure_built(self._declared_input_shape)

    gelu = staticmethod(_gelu_tanh)
    ln_eps = 1e-5

    def _embedding_out_dim(self, input_shape):
        if self._hidden:
            return self._hidden
        xs = [torch.zeros((2,) + tuple(s), dtype=torch.int64) for s in input_shape[:self.n_embedding_inputs]]
        with torch.no_grad():
            return self.embedding_layer(*xs).shape[-1]

    n_embedding_inputs = 2

    def build(self, input_shape):
        h = self._embedding_out_dim(input_shape)
        self.hidden_size = h
        inter = self.intermediate_size if self.intermediate_size > 0 else 4 * h
        self.blocks = nn.ModuleList([TransformerBlock(h, self.n_head, inter, self.hidden_drop, self.attn_drop, self.initializer_range,
                                                      not self.bidirectional, self.gelu, self.ln_eps) for _ in range(self.n_block)])
        self.pooler = FusedDense(h, h)

    def _mask(self, xs):
        return None

    def call(self, xs):
        e = self.embedding_layer(*xs[:self.n_embedding_inputs])
        mask = self._mask(xs)
        outs, h = [], e
        for b in self.blocks:
            h = b(h, mask)
            outs.append(h)
        pooled = torch.tanh(self.pooler(h[:, 0]).float())
        return outs + [pooled] if self.output_all_block else [h, pooled]

    def compute_output_shape(self, input_shape):
        L, h = self.seq_len, self.hidden_size
        return ([(L, h)] * self.n_block if self.output_all_block else [(L, h)]) + [(h,)]

    @classmethod
    def init(cls, vocab=40990, seq_len=77, n_block=12, hidden_drop=0.1, attn_drop=0.1, n_head=12, hidden_size=768,
             embedding_drop=0.1, initializer_range=0.02, bidirectional=False, output_all_block=False):
        if hidden_size < 0:
            raise TypeError("hidden_size must be greater than 0 with default embedding layer")
        emb = _SumEmbedding(vocab, hidden_size, initializer_range, embedding_drop)
        return cls(n_block, hidden_drop, attn_drop, n_head, initializer_range, bidirectional, output_all_block, emb,
                   input_shape=((seq_len,), (seq_len,)), hidden_size=hidden_size)


class _SumEmbedding(nn.Module):
    """Default TransformerLayer embedding: one table indexed by token ids and by position ids, summed
    (self_attention.py:212-226: positions are encoded as extra vocabulary entries)."""

    def __init__(self, vocab, hidden, init_range, drop):
        super().__init__()
        self.table = nn.Embedding(vocab, hidden)
        nn.init.normal_(self.table.weight, 0.0, init_range)
        self.drop = drop

    def forward(self, word, pos):
        e = self.table(word.long()) + self.table(pos.long())
        return F.dropout(e, self.drop, self.training)


class _BertEmbedding(nn.Module):
    def __init__(self, vocab, hidden, seq_len, init_range, drop):
        super().__init__()
        self.word, self.pos, self.tok = nn.Embedding(vocab, hidden), nn.Embedding(seq_len, hidden), nn.Embedding(2, hidden)
        for t in (self.word, self.pos, self.tok):
            nn.init.normal_(t.weight, 0.0, init_range)
        self.ln = nn.LayerNorm(hidden, eps=1e-12)
        self.drop = drop

    def forward(self, word, token_type, position):
        e = self.word(word.long()) + self.pos(position.long()) + self.tok(token_type.long())
        return F.dropout(self.ln(e), self.drop, self.training)


class BERT(TransformerLayer):
    """Inputs [token ids, token type ids, position ids, attention mask (B,1,1,L) of 1/0] ->
    [block outputs..., pooled] (`output_all_block=True` by default, as in the reference)."""
    gelu = staticmethod(_gelu_erf)
    ln_eps = 1e-12
    n_embedding_inputs = 3

    def __init__(self, n_block, n_head, intermediate_size, hidden_drop, attn_drop, initializer_range, output_all_block,
                 embedding_layer, input_shape, hidden_size=None, **kw):
        if len(input_shape) != 4:
            raise TypeError("BERT input must be a list of 4 ndarray (consisting of input sequence, sequence positions, "
                            "segment id, attention mask)")
        super().__init__(n_block, hidden_drop, attn_drop, n_head, initializer_range, True, output_all_block, embedding_layer,
                         input_shape, intermediate_size, hidden_size, **kw)

    def _mask(self, xs):
        return (1.0 - xs[3].float()) * -10000.0

    def compute_output_shape(self, input_shape):
        return super().compute_output_shape(input_shape)

    def ensure_built(self, input_shape):
        if not self.built and input_shape is not None:
            self.build(input_shape)
            self.built, self.input_shape_ = True, input_shape
            self.output_shape_ = self.compute_output_shape(input_shape)
        return self.output_shape_

    @classmethod
    def init(cls, vocab=40990, hidden_size=768, n_block=12, n_head=12, seq_len=512, intermediate_size=3072, hidden_drop=0.1,
             attn_drop=0.1, initializer_range=0.02, output_all_block=True, bigdl_type="float"):
        emb = _BertEmbedding(vocab, hidden_size, seq_len, initializer_range, hidden_drop)
        return cls(n_block, n_head, intermediate_size, hidden_drop, attn_drop, initializer_range, output_all_block, emb,
                   input_shape=((seq_len,), (seq_len,), (seq_len,), (1, 1, seq_len)), hidden_size=hidden_size)

    @staticmethod
    def init_from_existing_model(path, weight_path=None, input_seq_len=-1.0, hidden_drop=-1.0, attn_drop=-1.0,
                                 output_all_block=True, bigdl_type="float"):
        m = torch.load(path, map_location="cpu", weights_only=False)
        if not isinstance(m, BERT):
            raise TypeError(f"{path} does not hold a BERT layer saved by zoo-b200 (BigDL protobuf models cannot be read here)")
        if hidden_drop >= 0:
            m.hidden_drop = hidden_drop
        if attn_drop >= 0:
            m.attn_drop = attn_drop
        for b in m.blocks:
            b.hidden_drop, b.attn_drop = m.hidden_drop, m.attn_drop
        m.output_all_block = output_all_block
        return m


def layer_norm(x, w, b, e=1e-5):
    """Variable-level layer norm over the last dimension (self_attention.py `layer_norm`)."""
    from ... import autograd as A

    u = A.mean(x, axis=len(x.get_output_shape()) - 1, keepDims=True)
    s = A.mean(A.square(x - u), axis=len(x.get_output_shape()) - 1, keepDims=True)
    return (x - u) / A.sqrt(s + e) * w + b
