"""XShards / DataShards: a partitioned collection pre-processed in parallel and fed shard-wise to GPUs.

Reference: pyzoo/zoo/xshard/shard.py:20-100 (`DataShards.apply/collect`, `RayDataShards.repartition /
get_partitions`, `RayPartition`) where every shard is a Ray actor holding one pandas DataFrame.
Here shards are plain in-process objects (pandas DataFrames, dicts of ndarrays, anything);
`apply` / `transform_shard` map a function over them on the context's worker pool (or a thread pool --
pandas and numpy release the GIL for the heavy parts); in a multi-process run every rank keeps the
partitions `i % world == rank` (`collect_local`), so partition i lands on GPU `i % world` without a
shuffle -- the counterpart of `coalesce(nodeNumber)` (FeatureSet.scala:677-678).  The later Orca names
(`XShards.partition`, `transform_shard`, `num_partitions`, `to_torch_dataset`) are provided next to the
snapshot's own (`apply`, `collect`, `repartition`, `get_partitions`).
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor

import numpy as np


class DataShards:
    def apply(self, func, *args):  # pragma: no cover - interface
        raise NotImplementedError

    def collect(self):  # pragma: no cover - interface
        raise NotImplementedError


class RayPartition:
    def __init__(self, shard_list):
        self.shard_list = list(shard_list)

    def get_data(self):
        return list(self.shard_list)


def _chunks(seq, n):
    """Split `seq` into `n` contiguous runs of (almost) equal length."""
    k, m = divmod(len(seq), n)
    out, s = [], 0
    for i in range(n):
        e = s + k + (1 if i < m else 0)
        out.append(seq[s:e])
        s = e
    return out


class XShards(DataShards):
    def __init__(self, shards, context=None):
        self.shard_list = list(shards)
        self.partitions = [RayPartition([s]) for s in self.shard_list]
        self.context = context

    # ---- constructors ----------------------------------------------------------------------
    @classmethod
    def partition(cls, data, num_shards=None, context=None):
        """Split ndarrays / (nested) dicts, lists, tuples of ndarrays along axis 0 into `num_shards` shards."""
        import torch

        if num_shards is None:
            num_shards = max(1, torch.cuda.device_count()) if torch.cuda.is_available() else 4
        from ..util import nest

        flat = nest.flatten(data)
        n = len(flat[0])
        if any(len(a) != n for a in flat):
            raise ValueError("all arrays must have the same first dimension")
        num_shards = max(1, min(num_shards, n))
        bounds = np.linspace(0, n, num_shards + 1).astype(int)
        shards = [nest.pack_sequence_as(data, [np.asarray(a)[bounds[i]:bounds[i + 1]] for a in flat]) for i in range(num_shards)]
        return cls(shards, context)

    # ---- transformations -----------------------------------------------------------------------
    def _map(self, fn):
        ctx = self.context
        if ctx is not None and hasattr(ctx, "remote") and not getattr(ctx, "local_mode", True):
            return [r.get() for r in [ctx.remote(fn, s) for s in self.shard_list]]
        if len(self.shard_list) <= 1:
            return [fn(s) for s in self.shard_list]
        with ThreadPoolExecutor(min(len(self.shard_list), 16)) as ex:
            return list(ex.map(fn, self.shard_list))

    def transform_shard(self, func, *args):
        """Returns a NEW XShards of `func(shard, *args)`."""
        return XShards(self._map(lambda s: func(s, *args)), self.context)

    def apply(self, func, *args):
        """Reference semantics: in place, returns self (shard.py:50-63)."""
        self.shard_list = self._map(lambda s: func(s, *args))
        self.repartition(len(self.partitions))
        return self

    def collect(self):
        return list(self.shard_list)

    def collect_local(self, rank=None, world=None):
        """This rank's shards (shard i belongs to rank i % world)."""
        if rank is None:
            from ..common.nncontext import get_context

            c = get_context()
            rank, world = c.rank, c.world_size
        return [s for i, s in enumerate(self.shard_list) if i % world == rank]

    def num_partitions(self):
        return len(self.partitions)

    def repartition(self, num_partitions):
        num_partitions = max(1, min(int(num_partitions), max(1, len(self.shard_list))))
        self.partitions = [RayPartition(c) for c in _chunks(self.shard_list, num_partitions)]
        return self

    def get_partitions(self):
        return self.partitions

    def __len__(self):
        return sum(len(s) if hasattr(s, "__len__") and not isinstance(s, dict) else len(next(iter(s.values()))) for s in self.shard_list)

    def split(self, n):
        """Merge/split so that there are exactly `n` shards (pandas or dict-of-ndarray shards)."""
        import pandas as pd

        if not self.shard_list:
            return self
        if isinstance(self.shard_list[0], pd.DataFrame):
            df = pd.concat(self.shard_list, ignore_index=True)
            b = np.linspace(0, len(df), n + 1).astype(int)
            return XShards([df.iloc[b[i]:b[i + 1]].reset_index(drop=True) for i in range(n)], self.context)
        from ..util import nest

        flat = [nest.flatten(s) for s in self.shard_list]
        cat = [np.concatenate([f[k] for f in flat]) for k in range(len(flat[0]))]
        return XShards.partition(nest.pack_sequence_as(self.shard_list[0], cat), n, self.context)

    def to_torch_dataset(self, feature_cols=None, label_cols=None):
        import torch

        from ..feature.common import FeatureSet

        fs = FeatureSet.xshards(self, feature_cols, label_cols)
        ts = list(fs.features) + list(fs.labels or ())
        return torch.utils.data.TensorDataset(*ts)

    def to_feature_set(self, feature_cols=None, label_cols=None, **kw):
        from ..feature.common import FeatureSet

        return FeatureSet.xshards(self, feature_cols, label_cols, **kw)


RayDataShards = XShards
