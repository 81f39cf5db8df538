"""Cluster Serving: pub/sub batched inference (reference: pyzoo/zoo/serving, zoo/.../serving/*.scala)."""
from .client import API, InputQueue, OutputQueue  # noqa: F401
from .server import ClusterServing, ServingConfig  # noqa: F401
