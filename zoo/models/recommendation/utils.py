"""Feature-engineering helpers for the recommenders.

Reference: pyzoo/zoo/models/recommendation/utils.py:25-166 and
zoo/src/main/scala/.../models/recommendation/Utils.scala:38-327 (`hash_bucket`,
`categorical_from_vocab_list`, `get_boundaries`, `get_negative_samples`, `get_wide_tensor`,
`get_deep_tensors`, `row_to_sample`, `to_user_item_feature`, `prePadding`).  Rows are dicts /
pandas rows; the wide tensor is returned as the index list of the sparse multi-hot vector
(what `zoo.models.recommendation.WideAndDeep` consumes) instead of a BigDL sparse JTensor.
"""
from __future__ import annotations

import zlib

import numpy as np

from ...feature.common import Sample

__all__ = ["hash_bucket", "categorical_from_vocab_list", "get_boundaries", "get_negative_samples", "get_wide_tensor",
           "get_deep_tensors", "row_to_sample", "to_user_item_feature", "pre_padding", "bucketized_column", "cross_columns"]


def hash_bucket(content, bucket_size=1000, start=0):
    """Stable string hash bucket (Python's `hash` is salted per process; crc32 is not)."""
    return (zlib.crc32(str(content).encode()) % bucket_size + bucket_size) % bucket_size + start


def categorical_from_vocab_list(sth, vocab_list, default=-1, start=0):
    return vocab_list.index(sth) + start if sth in vocab_list else default + start


def get_boundaries(target, boundaries, default=-1, start=0):
    if target == "?":
        return default + start
    for i, b in enumerate(boundaries):
        if target < b:
            return i + start
    return len(boundaries) + start


bucketized_column = get_boundaries


def cross_columns(row, cols, bucket_size):
    """Hash of the '_'-joined values (Utils.scala `buckBuckets`)."""
    return hash_bucket("_".join(str(row[c]) for c in cols), bucket_size)


def get_negative_samples(indexed, neg_per_pos=1, seed=42):
    """For every positive (userId, itemId) row generate random unseen items with label 1; positives
    keep label 2 (Utils.scala:38-78: `getNegativeSamples`).  `indexed`: pandas DataFrame."""
    import pandas as pd

    rng = np.random.RandomState(seed)
    items = indexed["itemId"].unique()
    seen = set(zip(indexed["userId"].tolist(), indexed["itemId"].tolist()))
    rows = []
    for u in indexed["userId"].tolist():
        for _ in range(neg_per_pos):
            for _try in range(20):
                it = int(items[rng.randint(len(items))])
                if (u, it) not in seen:
                    rows.append((u, it, 1))
                    break
    return pd.DataFrame(rows, columns=["userId", "itemId", "label"])


def get_wide_tensor(row, column_info):
    """Indices (into the concatenated wide vector) of the active wide columns of `row`."""
    ci = column_info
    cols, dims = ci.wide_base_cols + ci.wide_cross_cols, ci.wide_base_dims + ci.wide_cross_dims
    idx, acc = [], 0
    for c, d in zip(cols, dims):
        v = int(row[c])
        idx.append(acc + v if 0 <= v < d else -1)
        acc += d
    return np.asarray(idx, dtype=np.int64)


def get_deep_tensors(row, column_info):
    """[indicator multi-hot, embedding ids, continuous values] (absent parts omitted)."""
    ci, out = column_info, []
    if ci.indicator_cols:
        ind, acc = np.zeros(sum(ci.indicator_dims), dtype=np.float32), 0
        for c, d in zip(ci.indicator_cols, ci.indicator_dims):
            v = int(row[c])
            if 0 <= v < d:
                ind[acc + v] = 1.0
            acc += d
        out.append(ind)
    if ci.embed_cols:
        out.append(np.asarray([int(row[c]) for c in ci.embed_cols], dtype=np.int64))
    if ci.continuous_cols:
        out.append(np.asarray([float(row[c]) for c in ci.continuous_cols], dtype=np.float32))
    return out


def row_to_sample(row, column_info, model_type="wide_n_deep"):
    label = np.asarray([int(row[column_info.label])], dtype=np.int64)
    if model_type == "wide":
        feats = [get_wide_tensor(row, column_info)]
    elif model_type == "deep":
        feats = get_deep_tensors(row, column_info)
    elif model_type == "wide_n_deep":
        feats = [get_wide_tensor(row, column_info)] + get_deep_tensors(row, column_info)
    else:
        raise TypeError(f"Unsupported model_type: {model_type}")
    return Sample(feats, label)


def to_user_item_feature(row, column_info, model_type="wide_n_deep"):
    from .recommender import UserItemFeature

    return UserItemFeature(row["userId"], row["itemId"], row_to_sample(row, column_info, model_type))


def pre_padding(seq, length, value=0):
    """Left-pad / left-truncate a sequence to `length` (Utils.scala:295-305 `prePadding`)."""
    seq = list(seq)[-length:]
    return [value] * (length - len(seq)) + seq
