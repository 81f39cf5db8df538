"""Trainer configuration for `MXNetTrainer` (the reference's counterpart is pyzoo/zoo/ray/mxnet/utils.py:21-46).

The configuration is the dict handed to every creator function (`model_creator(config)`, `data_creator(config, kv)`, ...).  Its
keys are part of the public contract -- user creator functions read `config["batch_size"]`, `config["optimizer"]`,
`config["optimizer_params"]`, `config["log_interval"]` and optionally `config["seed"]` -- so the key names are kept; building and
validating it is done here through a small schema instead of ad-hoc dict code.
"""
from __future__ import annotations

from ...util.launcher import free_port as find_free_port  # noqa: F401  (one port helper for the whole package)

_SCHEMA = {
    # key: (type check, default factory)
    "batch_size": (lambda v: isinstance(v, int) and v > 0, lambda: 32),
    "optimizer": (lambda v: isinstance(v, str) or callable(v), lambda: "sgd"),
    "optimizer_params": (lambda v: isinstance(v, dict), lambda: {"learning_rate": 0.01}),
    "log_interval": (lambda v: isinstance(v, int) and v > 0, lambda: 10),
}


def create_trainer_config(batch_size=None, optimizer=None, optimizer_params=None, log_interval=None, seed=None, extra_config=None):
    """Returns the config dict; unset fields take the schema defaults, `extra_config` entries are merged last and may override."""
    given = {"batch_size": batch_size, "optimizer": optimizer, "optimizer_params": optimizer_params or None, "log_interval": log_interval}
    cfg = {}
    for key, (valid, default) in _SCHEMA.items():
        v = given[key]
        if v is None:
            v = default()
        elif not valid(v):
            raise ValueError(f"trainer config: bad value for {key!r}: {v!r}")
        cfg[key] = v
    if seed is not None and seed is not False and seed != 0:
        cfg["seed"] = int(seed)
    if extra_config is not None:
        if not isinstance(extra_config, dict):
            raise TypeError("extra_config must be a dict")
        cfg.update(extra_config)
    return cfg
