"""Embedding layers (reference: layers/embeddings.py; Scala Embedding.scala:59-112, WordEmbedding.scala,
SparseEmbedding.scala:59-89).  On the GPU the lookups run on the embedding gather kernel K2 with the sparse
optimizer K4 in backward (zoo.ops.embedding)."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .....ops.embedding import EmbeddingTable, LookupSpec, multi_lookup
from .base import ZooKerasLayer


class Embedding(ZooKerasLayer):
    """Integer ids (batch, length) -> (batch, length, output_dim).  `zero_based_id=False` means ids start at 1
    (the reference adds the constant itself, Embedding.scala:82-101); `mask_zero` keeps row 0 at zero."""
    input_dtype = torch.int64

    def __init__(self, input_dim, output_dim, init="uniform", weights=None, trainable=True, input_length=None, W_regularizer=None,
                 input_shape=None, mask_zero=False, padding_value=0, zero_based_id=True, **kw):
        if input_length is not None and input_shape is None:
            input_shape = (int(input_length),)
        super().__init__(input_shape, **kw)
        self.input_dim, self.output_dim, self.init = int(input_dim), int(output_dim), init
        self.weights_init, self.trainable, self.mask_zero, self.zero_based_id = weights, trainable, mask_zero, zero_based_id

    def build(self, input_shape):
        self.table = EmbeddingTable(self.input_dim, self.output_dim, init=self.init if self.init in ("uniform", "normal", "zeros") else "xavier",
                                    sparse=False, padding_idx=0 if self.mask_zero else None)
        if self.weights_init is not None:
            with torch.no_grad():
                self.table.weight.copy_(torch.as_tensor(np.asarray(self.weights_init), dtype=torch.float32))
        if not self.trainable:
            self.table.weight.requires_grad_(False)

    def call(self, x):
        ids = x.long()
        if not self.zero_based_id:
            ids = ids - 1
        lead = ids.shape
        flat = ids.reshape(-1, 1)
        if flat.is_cuda and self.table.weight.requires_grad is False and not self.table.sparse:
            out = F.embedding(flat.squeeze(1), self.table.weight)
        else:
            out = multi_lookup(flat, [LookupSpec(self.table, 0)])
        out = out.reshape(*lead, self.output_dim)
        if self.mask_zero:
            out = out * (ids != 0).unsqueeze(-1).to(out.dtype)
        return out


class WordEmbedding(ZooKerasLayer):
    """Frozen (or trainable) pre-trained word vectors loaded from a GloVe text file (WordEmbedding.scala)."""
    input_dtype = torch.int64

    def __init__(self, embedding_file, word_index=None, trainable=False, input_length=None, input_shape=None, **kw):
        if input_length is not None and input_shape is None:
            input_shape = (int(input_length),)
        super().__init__(input_shape, **kw)
        self.embedding_file, self.word_index, self.trainable = embedding_file, word_index, trainable

    @staticmethod
    def get_word_index(embedding_file):
        idx = {}
        with open(embedding_file, encoding="utf-8") as f:
            for i, line in enumerate(f):
                idx[line.split(" ", 1)[0]] = i + 1
        return idx

    def build(self, input_shape):
        vecs, dim = {}, None
        with open(self.embedding_file, encoding="utf-8") as f:
            for line in f:
                parts = line.rstrip().split(" ")
                if self.word_index is None or parts[0] in self.word_index:
                    vecs[parts[0]] = np.asarray(parts[1:], dtype=np.float32)
                    dim = len(parts) - 1
        index = self.word_index or {w: i + 1 for i, w in enumerate(vecs)}
        w = np.zeros((max(index.values()) + 1, dim), np.float32)        # row 0 = unknown / padding
        for word, i in index.items():
            if word in vecs:
                w[i] = vecs[word]
        self.weight = nn.Parameter(torch.from_numpy(w), requires_grad=bool(self.trainable))
        self.output_dim = dim

    def call(self, x):
        return F.embedding(x.long(), self.weight)


class SparseEmbedding(ZooKerasLayer):
    """Bag lookup of a sparse (batch, n) id matrix with combiner sum | mean | sqrtn and optional max_norm
    (SparseEmbedding.scala:59-89 -> LookupTableSparse).  Input: dense int ids where -1 (or 0 with
    zero_based_id=False) marks "no entry", or a torch sparse tensor whose values are the ids."""
    input_dtype = torch.int64

    def __init__(self, input_dim, output_dim, combiner="sum", max_norm=-1.0, init="uniform", W_regularizer=None, input_shape=None,
                 zero_based_id=True, **kw):
        super().__init__(input_shape, **kw)
        self.input_dim, self.output_dim, self.combiner, self.max_norm = int(input_dim), int(output_dim), combiner.lower(), float(max_norm)
        self.init, self.zero_based_id = init, zero_based_id
        if self.combiner not in ("sum", "mean", "sqrtn"):
            raise ValueError(f"Unsupported combiner: {combiner}")

    def build(self, input_shape):
        self.table = EmbeddingTable(self.input_dim, self.output_dim, init=self.init if self.init in ("uniform", "normal", "zeros") else "xavier", sparse=False)

    def call(self, x):
        if x.is_sparse:
            x = x.to_dense() - (x.to_dense() == 0).long() if self.zero_based_id else x.to_dense()
        ids = x.long()
        if not self.zero_based_id:
            ids = ids - 1                                     # 0 (padding) becomes -1 = no entry
        if self.max_norm > 0:
            with torch.no_grad():
                n = self.table.weight.norm(dim=1, keepdim=True).clamp_min(1e-12)
                self.table.weight.mul_(torch.clamp(self.max_norm / n, max=1.0))
        out = multi_lookup(ids, [LookupSpec(self.table, 0, bag=ids.shape[1])]).float()
        cnt = (ids >= 0).sum(1, keepdim=True).clamp_min(1).float()
        if self.combiner == "mean":
            out = out / cnt
        elif self.combiner == "sqrtn":
            out = out / cnt.sqrt()
        return out


def prepare_embedding(embedding_file, word_index=None, randomize_unknown=False, normalize=False):
    from .....models.textmatching.knrm import prepare_embedding as _p

    return _p(embedding_file, word_index, randomize_unknown, normalize)
