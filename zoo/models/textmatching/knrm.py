"""KNRM: kernel-pooling neural ranking model (reference: pyzoo/zoo/models/textmatching/knrm.py:32-140,
zoo/.../models/textmatching/KNRM.scala:60-105).  Input: query and doc token ids concatenated
`[batch, text1_length + text2_length]`; shared embedding -> cosine translation matrix (embeddings are
L2-normalised when loaded from a file) -> `kernel_num` RBF kernels (the last one an exact-match kernel at mu=1
with `exact_sigma`) -> sum over doc, log1p, sum over query -> Dense(1) (sigmoid for classification)."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from ...ops.dense import FusedDense
from .text_matcher import TextMatcher


def prepare_embedding(embedding_file, word_index=None, randomize_unknown=False, normalize=False):
    """GloVe text file -> [vocab+1, dim] matrix (row 0 = padding); unknown words random-uniform(-0.05, 0.05) or zero."""
    vecs, dim = {}, None
    with open(embedding_file, encoding="utf-8") as f:
        for line in f:
            p = line.rstrip().split(" ")
            if word_index is None or p[0] in word_index:
                vecs[p[0]] = np.asarray(p[1:], dtype=np.float32)
                dim = len(p) - 1
    index = word_index or {w: i + 1 for i, w in enumerate(vecs)}
    rng = np.random.RandomState(0)
    n_rows = max(index.values()) + 1
    w = rng.uniform(-0.05, 0.05, (n_rows, dim)).astype(np.float32) if randomize_unknown else np.zeros((n_rows, dim), np.float32)
    w[0] = 0
    for word, i in index.items():
        if word in vecs:
            w[i] = vecs[word]
    if normalize:
        w = w / np.maximum(np.linalg.norm(w, axis=1, keepdims=True), 1e-12)
    return w


class KNRM(TextMatcher):
    def __init__(self, text1_length, text2_length, embedding_file=None, word_index=None, train_embed=True, kernel_num=21, sigma=0.1,
                 exact_sigma=0.001, target_mode="ranking", bigdl_type="float", vocab_size=None, embed_size=None, embed_weights=None):
        if embed_weights is None and embedding_file is not None:
            embed_weights = prepare_embedding(embedding_file, word_index, randomize_unknown=True, normalize=True)
        if embed_weights is not None:
            vocab_size, embed_size = embed_weights.shape
        elif not (vocab_size and embed_size):
            raise ValueError("KNRM needs embedding_file, embed_weights, or vocab_size + embed_size")
        super().__init__(text1_length, vocab_size, embed_size, embed_weights, train_embed, target_mode)
        assert kernel_num > 1, "kernel_num must be an int larger than 1"
        self.text2_length, self.kernel_num, self.sigma, self.exact_sigma = int(text2_length), int(kernel_num), float(sigma), float(exact_sigma)
        self._record_config(text1_length=self.text1_length, text2_length=self.text2_length, embedding_file=None, train_embed=train_embed,
                            kernel_num=self.kernel_num, sigma=self.sigma, exact_sigma=self.exact_sigma, target_mode=target_mode,
                            vocab_size=self.vocab_size, embed_size=self.embed_size)
        self.embedding = nn.Embedding(self.vocab_size, self.embed_size)
        if embed_weights is not None:
            with torch.no_grad():
                self.embedding.weight.copy_(torch.as_tensor(embed_weights, dtype=torch.float32))
        else:
            nn.init.uniform_(self.embedding.weight, -0.05, 0.05)
        self.embedding.weight.requires_grad_(bool(train_embed))
        mus, sigmas = [], []
        for i in range(self.kernel_num):
            mu = 1.0 / (self.kernel_num - 1) + (2.0 * i) / (self.kernel_num - 1) - 1.0
            s = self.sigma
            if mu > 1.0:
                mu, s = 1.0, self.exact_sigma
            mus.append(mu)
            sigmas.append(s)
        self.register_buffer("mu", torch.tensor(mus).view(1, 1, 1, -1))
        self.register_buffer("sig", torch.tensor(sigmas).view(1, 1, 1, -1))
        self.out = FusedDense(self.kernel_num, 1, init="uniform")

    def forward_logits(self, x):
        e = self.embedding(x.long())
        q, d = e[:, :self.text1_length], e[:, self.text1_length:self.text1_length + self.text2_length]
        mm = torch.bmm(q, d.transpose(1, 2)).unsqueeze(-1)                      # [B, L1, L2, 1]
        k = torch.exp(-0.5 * (mm - self.mu) ** 2 / (self.sig * self.sig))      # [B, L1, L2, K]
        phi = torch.log1p(k.sum(2)).sum(1)                                     # [B, K]
        return self.out(phi).float()

    def forward(self, x):
        y = self.forward_logits(x)
        return torch.sigmoid(y) if self.target_mode == "classification" else y

    def logits_kind(self):
        return "sigmoid" if self.target_mode == "classification" else None
