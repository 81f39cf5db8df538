"""Kept for import compatibility: the dispatch lives in `zoo.ops.rnn.run`."""
from .rnn import run  # noqa: F401
