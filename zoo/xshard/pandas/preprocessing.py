"""`read_csv` / `read_json` (/ `read_parquet`) -> XShards of pandas DataFrames.

Reference: pyzoo/zoo/xshard/pandas/preprocessing.py:26-189 -- `file_path` is one file, several comma
separated files, or a directory; files are spread over the workers and every worker loads its files
into one DataFrame.  Local paths (and `file://`) are supported; hdfs:// and s3:// need clients that are
not on this box and raise a clear error.
"""
from __future__ import annotations

import os

from ..shard import XShards, _chunks


def list_files(file_path, file_type):
    paths = []
    for p in str(file_path).split(","):
        p = p.strip()
        if p.startswith("file://"):
            p = p[len("file://"):]
        if "://" in p:
            raise NotImplementedError(f"{p.split('://')[0]}:// paths need a remote file-system client that is not available here")
        if os.path.isdir(p):
            paths += sorted(os.path.join(p, f) for f in os.listdir(p) if f.endswith("." + file_type))
        elif os.path.exists(p):
            paths.append(p)
        else:
            raise FileNotFoundError(p)
    if not paths:
        raise Exception(f"no {file_type} files found under {file_path}")
    return paths


def _load(files, file_type, kwargs):
    import pandas as pd

    reader = {"csv": pd.read_csv, "json": lambda f, **k: pd.read_json(f, **({"orient": "columns", "lines": False} | k)),
              "parquet": pd.read_parquet}[file_type]
    dfs = [reader(f, **kwargs) for f in files]
    return dfs[0] if len(dfs) == 1 else pd.concat(dfs, ignore_index=True)


def read_file(context, file_path, file_type, **kwargs):
    files = list_files(file_path, file_type)
    n = getattr(context, "num_workers", None) or getattr(context, "world_size", None) or 1
    groups = [g for g in _chunks(files, max(1, min(n, len(files)))) if g]
    shards = XShards(groups, context).transform_shard(_load, file_type, kwargs)
    return shards


def read_csv(file_path, context=None, **kwargs):
    return read_file(context, file_path, "csv", **kwargs)


def read_json(file_path, context=None, **kwargs):
    return read_file(context, file_path, "json", **kwargs)


def read_parquet(file_path, context=None, **kwargs):
    return read_file(context, file_path, "parquet", **kwargs)


def read_file_ray(context, file_path, file_type, **kwargs):
    return read_file(context, file_path, file_type, **kwargs)


class RayPandasShard:
    """One shard holding a pandas DataFrame (an actor in the reference; a plain object here, or a `RayContext.actor`)."""

    def __init__(self, data=None):
        self.data = data

    def read_file_partitions(self, paths, file_type, **kwargs):
        self.data = _load(list(paths), file_type, kwargs)
        return 0

    def apply(self, func, *args):
        self.data = func(self.data, *args)
        return 0

    def get_data(self):
        return self.data
