"""NeuralCF: neural collaborative filtering (GMF + MLP).

Reference: pyzoo/zoo/models/recommendation/neuralcf.py:30-99 and
zoo/src/main/scala/.../models/recommendation/NeuralCF.scala:45-103.  Same constructor arguments
and the same graph: four embedding tables of `count + 1` rows (ids are 1-based; row 0 is the
reference's AddConstant offset slot), MLP tower over concat(user, item) with ReLU Dense layers,
GMF = element-wise product of the MF embeddings, concat -> Dense(class_num) -> softmax.
Python-side initialisation is `uniform` (neuralcf.py:73-87).

B200 execution: both MLP lookups are ONE gather into the tower's input buffer (concat fused
away), the GMF product is formed inside the gather kernel, the tower runs on the tcgen05 GEMM
kernel, the softmax lives in the fused loss kernel during training, and in sparse mode the four
tables are updated inside backward by the fused sparse optimizer (touched rows only).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ...ops.dense import FusedDense, FusedMLP
from ...ops.embedding import EmbeddingTable, LookupSpec, multi_lookup
from .recommender import Recommender


class NeuralCF(Recommender):
    def __init__(self, user_count, item_count, class_num, user_embed=20, item_embed=20, hidden_layers=(40, 20, 10),
                 include_mf=True, mf_embed=20, bigdl_type="float", sparse=True, world=1, rank=0, shadow_bf16=None,
                 device=None):
        super().__init__()
        self.user_count, self.item_count, self.class_num = int(user_count), int(item_count), int(class_num)
        self.user_embed, self.item_embed = int(user_embed), int(item_embed)
        self.hidden_layers = [int(u) for u in hidden_layers]
        self.include_mf, self.mf_embed = bool(include_mf), int(mf_embed)
        self._record_config(user_count=self.user_count, item_count=self.item_count, class_num=self.class_num,
                            user_embed=self.user_embed, item_embed=self.item_embed, hidden_layers=self.hidden_layers,
                            include_mf=self.include_mf, mf_embed=self.mf_embed)
        kw = dict(init="uniform", sparse=sparse, world=world, rank=rank, shadow_bf16=shadow_bf16, device=device)
        self.mlp_user_embed = EmbeddingTable(self.user_count + 1, self.user_embed, **kw)
        self.mlp_item_embed = EmbeddingTable(self.item_count + 1, self.item_embed, **kw)
        self.mlp = FusedMLP(self.user_embed + self.item_embed, self.hidden_layers, activation="relu")
        last_in = self.hidden_layers[-1]
        if self.include_mf:
            assert self.mf_embed > 0
            self.mf_user_embed = EmbeddingTable(self.user_count + 1, self.mf_embed, **kw)
            self.mf_item_embed = EmbeddingTable(self.item_count + 1, self.mf_embed, **kw)
            last_in += self.mf_embed
        # The last layer's dgrad applies the ReLU mask of the MLP output (first columns of its
        # input) in its epilogue, so the tower's last ReLU needs no separate backward pass.
        fuse_mask = self.hidden_layers[-1] % 8 == 0
        self.linear_last = FusedDense(last_in, self.class_num, activation=None, relu_input=fuse_mask,
                                      relu_input_cols=self.hidden_layers[-1] if self.include_mf else 0)
        self.mlp[-1].skip_relu_bwd = fuse_mask

    graph_safe = True  # static shapes, no host-side control flow: the step can be CUDA-graph captured

    def logits_kind(self):
        return "softmax"

    def forward_logits(self, x):
        """x: [B, 2] (user_id, item_id), 1-based like the reference."""
        return self.linear_last(self._head_input(x))

    def training_loss(self, x, y, criterion):
        """Loss of one training batch (the engine's fused route, `TrainingEngine.forward_loss`): the head -- Dense(class_num) +
        softmax + NLL and its whole backward -- runs as ONE kernel over the head's input when it applies (zoo.ops.head_loss),
        layer by layer otherwise."""
        from ...ops import head_loss as _hl

        hc = self._head_input(x)
        if _hl.applicable(hc, self.linear_last, criterion, y):
            return _hl.head_loss(hc, self.linear_last, criterion, y)
        return criterion.from_logits(self.linear_last(hc), y)

    def _head_input(self, x):
        """Everything below the last Dense layer: [tower output | GMF product]."""
        ids = x.long() if x.dtype != torch.int64 else x
        specs = [LookupSpec(self.mlp_user_embed, 0), LookupSpec(self.mlp_item_embed, 1)]
        if not self.include_mf:
            return self.mlp(multi_lookup(ids, specs))
        # ONE lookup for all four tables: [MLP-user | MLP-item | MF-user * MF-item]; the tower reads the first window, the
        # GMF product is the second.  Tables of one id column share their sparse update (zoo.ops.embedding_group).
        specs.append(LookupSpec(self.mf_user_embed, 0, mul_table=self.mf_item_embed, mul_col=1))
        hl = self.hidden_layers[-1]
        h_in, mf, wide = multi_lookup(ids, specs, splits=[self.user_embed + self.item_embed, self.mf_embed],
                                      embed=(1, hl + self.mf_embed, hl))
        if wide is not None and hl % 8 == 0:
            # fused concat: the gather wrote the GMF product into columns [hl, hl + mf) of `wide`, the last tower layer writes
            # its output into columns [0, hl)
            h = self.mlp(h_in, out=wide[:, :hl])          # three-layer ReLU towers run as one kernel (zoo.ops.dense.FusedMLP.forward)
            bits = getattr(h, "_zoo_relu_bits", None)
            from ...ops.embedding_group import join_columns

            hc = join_columns(wide, h, mf)
        else:
            h = self.mlp(h_in)
            bits = getattr(h, "_zoo_relu_bits", None)
            hc = torch.cat([h, mf.to(h.dtype)], dim=1)
        if bits is not None:                # sign bits of the MLP half travel with the concatenated activation
            hc._zoo_relu_bits = bits
        if self.mf_user_embed.world > 1 and hc.is_cuda and torch.is_grad_enabled():
            # sharded tables: the last layer writes its input gradient into peer-readable memory, so the owners of the MF rows
            # read their gradient rows in place (no staging copy of the MF half)
            from ...ops.embedding import ShardRuntime

            # (keyed by the table's serial number: an id() could be reused by a later object on one rank only, and symmetric
            # allocations are collective)
            hc._zoo_dx_out = ShardRuntime.get().stage(("ncf-last-dx", self.mf_user_embed.uid, hc.shape[0]), tuple(hc.shape), torch.bfloat16)[1]
        return hc

    def forward(self, x):
        return torch.softmax(self.forward_logits(x).float(), dim=-1)
