"""Orca names (the later public API of the same components, BASELINE.json / SURVEY 2.7)."""
from ..common.nncontext import init_orca_context, stop_orca_context  # noqa: F401


class OrcaContext:
    log_output = False
    pandas_read_backend = "pandas"
    serialize_data_creator = False
