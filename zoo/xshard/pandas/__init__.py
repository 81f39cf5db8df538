from .preprocessing import read_csv, read_json, read_parquet  # noqa: F401
