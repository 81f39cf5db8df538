"""SessionRecommender: GRU over the item-embedding sequence of a session (+ optional MLP over the
summed purchase history), softmax over items.

Reference: pyzoo/zoo/models/recommendation/session_recommender.py:28-140 and
zoo/src/main/scala/.../models/recommendation/SessionRecommender.scala:45-147
(`recommendForSession` returns the top-k (item, probability) per session).
The GRU stack runs on the persistent recurrent kernel K6 (`zoo.ops.rnn`) on GPU.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from ...ops.dense import FusedDense, FusedMLP
from ...ops.rnn import GRU
from .recommender import Recommender


class SessionRecommender(Recommender):
    def __init__(self, item_count, item_embed, rnn_hidden_layers=(40, 20), session_length=0, include_history=False,
                 mlp_hidden_layers=(40, 20), history_length=0, bigdl_type="float"):
        super().__init__()
        assert session_length > 0, "session_length should align with input features"
        if include_history:
            assert history_length > 0, "history_length should align with input features"
        self.item_count, self.item_embed = int(item_count), int(item_embed)
        self.rnn_hidden_layers = [int(u) for u in rnn_hidden_layers]
        self.mlp_hidden_layers = [int(u) for u in mlp_hidden_layers]
        self.include_history, self.session_length, self.history_length = bool(include_history), int(session_length), int(history_length)
        self._record_config(item_count=self.item_count, item_embed=self.item_embed, rnn_hidden_layers=self.rnn_hidden_layers,
                            session_length=self.session_length, include_history=self.include_history,
                            mlp_hidden_layers=self.mlp_hidden_layers, history_length=self.history_length)
        self.session_table = nn.Embedding(self.item_count + 1, self.item_embed)
        nn.init.uniform_(self.session_table.weight, -0.05, 0.05)
        grus, prev = [], self.item_embed
        for h in self.rnn_hidden_layers:
            grus.append(GRU(prev, h))
            prev = h
        self.grus = nn.ModuleList(grus)
        self.rnn_out = FusedDense(prev, self.item_count)
        if self.include_history:
            self.his_table = nn.Embedding(self.item_count + 1, self.item_embed)
            nn.init.uniform_(self.his_table.weight, -0.05, 0.05)
            self.mlp = FusedMLP(self.item_embed, self.mlp_hidden_layers, activation="relu")
            self.mlp_out = FusedDense(self.mlp_hidden_layers[-1], self.item_count)

    def logits_kind(self):
        return "softmax"

    def forward_logits(self, session, history=None):
        h = self.session_table(session.long())
        for g in self.grus[:-1]:
            h = g(h, return_sequences=True)
        h = self.grus[-1](h, return_sequences=False)
        out = self.rnn_out(h).float()
        if self.include_history:
            s = self.his_table(history.long()).sum(1)
            out = out + self.mlp_out(self.mlp(s)).float()
        return out

    def forward(self, session, history=None):
        return torch.softmax(self.forward_logits(session, history), dim=-1)

    def recommend_for_user(self, feature_rdd, max_items):
        raise Exception("recommend_for_user: Unsupported for SessionRecommender")

    def recommend_for_item(self, feature_rdd, max_users):
        raise Exception("recommend_for_item: Unsupported for SessionRecommender")

    def predict_user_item_pair(self, feature_rdd):
        raise Exception("predict_user_item_pair: Unsupported for SessionRecommender")

    def recommend_for_session(self, sessions, max_items, zero_based_label=True):
        """sessions: list of Samples / arrays (session[, history]).  Returns, per session, a list of
        (item, probability) of length `max_items`, best first."""
        if isinstance(sessions, tuple) and all(hasattr(s, "ndim") and s.ndim == 2 for s in sessions):
            xs = [np.asarray(s) for s in sessions]          # already batched: (session[B,L], history[B,H])
        else:
            feats = [(s.features if hasattr(s, "features") else (list(s) if isinstance(s, (list, tuple)) else [s])) for s in sessions]
            xs = [np.stack([np.asarray(f[i]) for f in feats]) for i in range(len(feats[0]))]
        probs = torch.as_tensor(np.asarray(self.predict(xs[0] if len(xs) == 1 else xs)))
        top = probs.topk(int(max_items), dim=-1)
        off = 0 if zero_based_label else 1
        return [[(int(i) + off, float(p)) for i, p in zip(idx, val)] for idx, val in zip(top.indices.tolist(), top.values.tolist())]
