"""WideAndDeep recommender.

Reference: pyzoo/zoo/models/recommendation/wide_and_deep.py:29-223 and
zoo/src/main/scala/.../models/recommendation/WideAndDeep.scala:54-347.

* wide part  = `SparseDense(class_num)` over a multi-hot sparse vector of `wide_dims` columns,
  i.e. a sum of rows of a `[wide_dims, class_num]` table -> the pooled (bag) mode of the
  embedding gather kernel K2;
* deep part  = indicator multi-hots (+) one Embedding per `embed_cols` (tables of `dim + 1` rows,
  `normal` init) (+) continuous values -> ReLU Dense tower -> Dense(class_num, relu)
  (the Python reference keeps a ReLU on the last deep layer, wide_and_deep.py:179);
* output     = softmax(wide + deep).

Inputs (tuple, in this order, absent parts omitted -- same order as the reference's
`Model([input_wide] + input_deep)`):
    wide_ids  int64 [B, n_wide]  global column indices into the wide vector, -1 = no entry
    ind       float [B, sum(indicator_dims)]
    emb_ids   int64 [B, len(embed_cols)]
    cont      float [B, len(continuous_cols)]
Sub-modules are named ``wide_part`` / ``deep_part`` so per-sub-module optimizers
(FTRL for wide, Adagrad for deep, CensusWideAndDeep.scala:122-132) can be assigned.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ...ops.dense import FusedDense, FusedMLP
from ...ops.embedding import EmbeddingTable, LookupSpec, multi_lookup
from .recommender import Recommender


class ColumnFeatureInfo:
    """Column description shared by the model and its feature generation (wide_and_deep.py:29-92)."""

    def __init__(self, wide_base_cols=None, wide_base_dims=None, wide_cross_cols=None, wide_cross_dims=None,
                 indicator_cols=None, indicator_dims=None, embed_cols=None, embed_in_dims=None, embed_out_dims=None,
                 continuous_cols=None, label="label", bigdl_type="float"):
        self.wide_base_cols = list(wide_base_cols or [])
        self.wide_base_dims = [int(d) for d in (wide_base_dims or [])]
        self.wide_cross_cols = list(wide_cross_cols or [])
        self.wide_cross_dims = [int(d) for d in (wide_cross_dims or [])]
        self.indicator_cols = list(indicator_cols or [])
        self.indicator_dims = [int(d) for d in (indicator_dims or [])]
        self.embed_cols = list(embed_cols or [])
        self.embed_in_dims = [int(d) for d in (embed_in_dims or [])]
        self.embed_out_dims = [int(d) for d in (embed_out_dims or [])]
        self.continuous_cols = list(continuous_cols or [])
        self.label = label

    def to_dict(self):
        return {k: v for k, v in self.__dict__.items()}

    def __reduce__(self):
        return ColumnFeatureInfo, (self.wide_base_cols, self.wide_base_dims, self.wide_cross_cols, self.wide_cross_dims,
                                   self.indicator_cols, self.indicator_dims, self.embed_cols, self.embed_in_dims,
                                   self.embed_out_dims, self.continuous_cols, self.label)

    def __str__(self):
        return "ColumnFeatureInfo " + str(self.to_dict())


class _WidePart(nn.Module):
    def __init__(self, wide_dims, class_num, n_cols, sparse):
        super().__init__()
        self.n_cols = n_cols
        self.table = EmbeddingTable(wide_dims, class_num, init="zeros", sparse=sparse)
        self.bias = nn.Parameter(torch.zeros(class_num))

    def forward(self, wide_ids):
        out = multi_lookup(wide_ids, [LookupSpec(self.table, 0, bag=wide_ids.shape[1])])
        return out.float() + self.bias


class _DeepPart(nn.Module):
    def __init__(self, info_dims, hidden_layers, class_num, sparse):
        super().__init__()
        ind_dims, embed_in, embed_out, n_cont = info_dims
        self.n_ind, self.n_cont = sum(ind_dims), n_cont
        self.embeds = nn.ModuleList([EmbeddingTable(i + 1, o, init="normal", sparse=sparse) for i, o in zip(embed_in, embed_out)])
        width = self.n_ind + sum(embed_out) + n_cont
        if width == 0:
            raise TypeError("Empty deep model")
        self.width = width
        self.mlp = FusedMLP(width, list(hidden_layers), activation="relu")
        self.last = FusedDense(hidden_layers[-1], class_num, activation="relu", relu_input=True)
        self.mlp[-1].skip_relu_bwd = True

    def forward(self, ind, emb_ids, cont):
        parts = []
        if ind is not None:
            parts.append(ind)
        if emb_ids is not None and len(self.embeds):
            e = multi_lookup(emb_ids, [LookupSpec(t, i) for i, t in enumerate(self.embeds)])
            parts.append(e)
        if cont is not None:
            parts.append(cont)
        dt = torch.bfloat16 if parts[0].is_cuda else torch.float32
        h = torch.cat([p.to(dt) for p in parts], dim=1) if len(parts) > 1 else parts[0].to(dt)
        return self.last(self.mlp(h)).float()


class WideAndDeep(Recommender):
    def __init__(self, class_num, column_info, model_type="wide_n_deep", hidden_layers=(40, 20, 10), bigdl_type="float",
                 sparse=True):
        super().__init__()
        if isinstance(column_info, dict):
            column_info = ColumnFeatureInfo(**{k: v for k, v in column_info.items() if k != "bigdl_type"})
        ci = column_info
        assert len(ci.wide_base_cols) == len(ci.wide_base_dims), "size of wide_base_columns should match"
        assert len(ci.wide_cross_cols) == len(ci.wide_cross_dims), "size of wide_cross_columns should match"
        assert len(ci.indicator_cols) == len(ci.indicator_dims), "size of wide_indicator_columns should match"
        assert len(ci.embed_cols) == len(ci.embed_in_dims) == len(ci.embed_out_dims), "size of embed columns should match"
        if model_type not in ("wide", "deep", "wide_n_deep"):
            raise TypeError(f"Unsupported model_type: {model_type}")
        self.class_num, self.model_type, self.column_info = int(class_num), model_type, ci
        self.hidden_layers = [int(u) for u in hidden_layers]
        self._record_config(class_num=self.class_num, column_info=ci.to_dict(), model_type=model_type,
                            hidden_layers=self.hidden_layers)
        self.wide_dims = sum(ci.wide_base_dims) + sum(ci.wide_cross_dims)
        self.has_ind, self.has_emb, self.has_cont = bool(ci.indicator_dims), bool(ci.embed_in_dims), bool(ci.continuous_cols)
        if model_type in ("wide", "wide_n_deep"):
            self.wide_part = _WidePart(self.wide_dims, self.class_num, len(ci.wide_base_dims) + len(ci.wide_cross_dims), sparse)
        if model_type in ("deep", "wide_n_deep"):
            self.deep_part = _DeepPart((ci.indicator_dims, ci.embed_in_dims, ci.embed_out_dims, len(ci.continuous_cols)),
                                       self.hidden_layers, self.class_num, sparse)

    graph_safe = True  # static shapes, no host-side control flow: the step can be CUDA-graph captured

    def logits_kind(self):
        return "softmax"

    def _split(self, xs):
        xs = list(xs)
        wide = xs.pop(0) if self.model_type in ("wide", "wide_n_deep") else None
        ind = xs.pop(0) if (self.model_type != "wide" and self.has_ind) else None
        emb = xs.pop(0) if (self.model_type != "wide" and self.has_emb) else None
        cont = xs.pop(0) if (self.model_type != "wide" and self.has_cont) else None
        return wide, ind, emb, cont

    def forward_logits(self, *xs):
        wide, ind, emb, cont = self._split(xs)
        out = None
        if wide is not None:
            out = self.wide_part(wide.long())
        if self.model_type != "wide":
            d = self.deep_part(ind, emb.long() if emb is not None else None, cont)
            out = d if out is None else out + d
        return out

    def forward(self, *xs):
        if len(xs) == 1 and isinstance(xs[0], (tuple, list)):
            xs = tuple(xs[0])
        return torch.softmax(self.forward_logits(*xs).float(), dim=-1)

    def example_inputs(self, n, rng=None, device=None):
        """Random inputs of the right shapes and id ranges (tests, benchmarks, tracing)."""
        import numpy as np

        rng = rng or np.random.RandomState(0)
        ci, out = self.column_info, []
        if self.model_type in ("wide", "wide_n_deep"):
            dims = ci.wide_base_dims + ci.wide_cross_dims
            offs = np.concatenate([[0], np.cumsum(dims)[:-1]])
            out.append(torch.from_numpy(np.stack([rng.randint(0, d, n) + o for d, o in zip(dims, offs)], 1).astype(np.int64)))
        if self.model_type != "wide":
            if self.has_ind:
                ind = np.zeros((n, sum(ci.indicator_dims)), np.float32)
                o = 0
                for d in ci.indicator_dims:
                    ind[np.arange(n), o + rng.randint(0, d, n)] = 1.0
                    o += d
                out.append(torch.from_numpy(ind))
            if self.has_emb:
                out.append(torch.from_numpy(np.stack([rng.randint(1, d + 1, n) for d in ci.embed_in_dims], 1).astype(np.int64)))
            if self.has_cont:
                out.append(torch.from_numpy(rng.rand(n, len(ci.continuous_cols)).astype(np.float32)))
        if device is not None:
            out = [t.to(device) for t in out]
        return tuple(out)
