from .shard import DataShards, RayDataShards, RayPartition, XShards  # noqa: F401
