"""Helpers of the MXNet trainer (reference: pyzoo/zoo/ray/mxnet/utils.py:21-46)."""
import socket
from contextlib import closing


def find_free_port():
    with closing(socket.socket(socket.AF_INET, socket.SOCK_STREAM)) as s:
        s.bind(("", 0))
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        return s.getsockname()[1]


def create_trainer_config(batch_size=32, optimizer="sgd", optimizer_params=None, log_interval=10, seed=None, extra_config=None):
    if not optimizer_params:
        optimizer_params = {"learning_rate": 0.01}
    config = {"batch_size": batch_size, "optimizer": optimizer, "optimizer_params": optimizer_params, "log_interval": log_interval}
    if seed:
        config["seed"] = seed
    if extra_config:
        assert isinstance(extra_config, dict), "extra_config must be a dict"
        config.update(extra_config)
    return config
