"""`read_csv` / `read_json` (/ `read_parquet`) -> XShards of pandas DataFrames.

Reference: pyzoo/zoo/xshard/pandas/preprocessing.py:26-189 -- `file_path` is one file, several comma
separated files, or a directory; files are spread over the workers and every worker loads its files
into one DataFrame.  Local paths, `file://`, and every scheme `zoo.util.filesystem` resolves (hdfs://, s3://, registered
schemes) are supported.
"""
from __future__ import annotations

import os

from ..shard import XShards, _chunks


def list_files(file_path, file_type):
    """One file, several comma separated files, or directories -- local, hdfs:// or s3:// (zoo.util.filesystem)."""
    from ...util import filesystem as zfs

    paths = []
    for p in str(file_path).split(","):
        p = p.strip()
        if not zfs.exists(p):
            raise FileNotFoundError(p)
        paths += zfs.ls(p, suffix="." + file_type) if zfs.isdir(p) else [zfs.local_path(p) if zfs.is_local(p) else p]
    if not paths:
        raise Exception(f"no {file_type} files found under {file_path}")
    return paths


def _load(files, file_type, kwargs):
    import pandas as pd

    from ...util import filesystem as zfs

    def read_json_file(f, **k):
        """JSON lines, as the reference reads them (`orient="columns", lines=True`); a file that is one JSON document is read as such."""
        if "lines" in k:
            return pd.read_json(f, **({"orient": "columns"} | k))
        data = f.read() if hasattr(f, "read") else open(f, "rb").read()
        import io

        order = (False, True) if b"\n" not in data.strip() else (True, False)      # a single line is one document, not one record
        try:
            return pd.read_json(io.BytesIO(data), **({"orient": "columns", "lines": order[0]} | k))
        except ValueError:
            return pd.read_json(io.BytesIO(data), **({"orient": "columns", "lines": order[1]} | k))

    reader = {"csv": pd.read_csv, "json": read_json_file, "parquet": pd.read_parquet}[file_type]

    def read_one(f):
        if zfs.is_local(f):
            return reader(zfs.local_path(f), **kwargs)
        with zfs.open_file(f, "rb") as fh:          # remote: stream through pyarrow
            import io

            return reader(io.BytesIO(fh.read()), **kwargs)

    dfs = [read_one(f) for f in files]
    return dfs[0] if len(dfs) == 1 else pd.concat(dfs, ignore_index=True)


def read_file(context, file_path, file_type, **kwargs):
    files = list_files(file_path, file_type)
    # one DataFrame per worker of the context (the reference spreads the files over its Ray nodes); without a context, one per file
    n = getattr(context, "num_workers", None) or getattr(context, "world_size", None) or len(files)
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
