"""Recommender base class: pair prediction and top-N recommendation.

Reference: pyzoo/zoo/models/recommendation/recommender.py:29-135 and
zoo/src/main/scala/.../models/recommendation/Recommender.scala:36-105:
`predictUserItemPair` (arg-max class + its probability per pair), `recommendForUser` /
`recommendForItem` (group by user/item, sort by (-prediction, -probability), take N).
The Spark groupByKey + sort becomes a segmented top-N (`zoo.ops.recommend`, kernel K15 on GPU).
"""
from __future__ import annotations

import numpy as np
import torch

from ..common.zoo_model import KerasZooModel


class UserItemFeature:
    """(user_id, item_id, sample): `sample` is a Sample / ndarray / tuple of ndarrays holding the
    model input of this pair."""

    def __init__(self, user_id, item_id, sample, bigdl_type="float"):
        self.user_id, self.item_id, self.sample = int(user_id), int(item_id), sample

    def __reduce__(self):
        return UserItemFeature, (self.user_id, self.item_id, self.sample)

    def __str__(self):
        return f"UserItemFeature [user_id: {self.user_id}, item_id: {self.item_id}, {self.sample}]"


class UserItemPrediction:
    def __init__(self, user_id, item_id, prediction, probability, bigdl_type="float"):
        self.user_id, self.item_id, self.prediction, self.probability = int(user_id), int(item_id), int(prediction), float(probability)

    def __reduce__(self):
        return UserItemPrediction, (self.user_id, self.item_id, self.prediction, self.probability)

    def __str__(self):
        return (f"UserItemPrediction [user_id: {self.user_id}, item_id: {self.item_id}, "
                f"prediction: {self.prediction}, probability: {self.probability}]")

    __repr__ = __str__


def _features_of(sample):
    if hasattr(sample, "features"):
        return [np.asarray(f) for f in sample.features]
    if isinstance(sample, (list, tuple)):
        return [np.asarray(f) for f in sample]
    return [np.asarray(sample)]


class Recommender(KerasZooModel):
    def _pairs_to_arrays(self, feature_rdd):
        feats = list(feature_rdd)
        users = np.asarray([f.user_id for f in feats], dtype=np.int64)
        items = np.asarray([f.item_id for f in feats], dtype=np.int64)
        cols = [_features_of(f.sample) for f in feats]
        xs = [np.stack([c[i] for c in cols]) for i in range(len(cols[0]))] if cols else []
        return users, items, (xs[0] if len(xs) == 1 else xs)

    def predict_user_item_pair(self, feature_rdd, batch_size=65536):
        """list of UserItemFeature -> list of UserItemPrediction (1-based predicted class, its probability)."""
        users, items, x = self._pairs_to_arrays(feature_rdd)
        if len(users) == 0:
            return []
        probs = np.asarray(self.predict(x, batch_size=batch_size))
        cls_ = probs.argmax(-1)
        p = probs[np.arange(len(cls_)), cls_]
        return [UserItemPrediction(u, i, c + 1, q) for u, i, c, q in zip(users, items, cls_, p)]

    def _recommend(self, feature_rdd, n, by_user):
        from ...ops.recommend import segmented_topn

        preds = self.predict_user_item_pair(feature_rdd)
        if not preds:
            return []
        keys = np.asarray([(p.user_id if by_user else p.item_id) for p in preds], dtype=np.int64)
        cls_ = np.asarray([p.prediction for p in preds], dtype=np.int64)
        prob = np.asarray([p.probability for p in preds], dtype=np.float32)
        keep = segmented_topn(keys, cls_, prob, int(n))
        return [preds[i] for i in keep]

    def recommend_for_user(self, feature_rdd, max_items):
        return self._recommend(feature_rdd, max_items, True)

    def recommend_for_item(self, feature_rdd, max_users):
        return self._recommend(feature_rdd, max_users, False)
