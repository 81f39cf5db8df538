from ...xshard.shard import XShards  # noqa: F401
from . import pandas  # noqa: F401
