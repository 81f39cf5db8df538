from .mxnet_trainer import MXNetTrainer  # noqa: F401
