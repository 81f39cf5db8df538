"""Helpers of the net package (reference: pyzoo/zoo/pipeline/api/net/utils.py:27-160).  `find_tensors` / `find_placeholders`
walk a TensorFlow graph (not available here); `to_bigdl_optim_method` turns an optimizer description into the training
engine's optimizer configuration."""
from __future__ import annotations

import torch


def _needs_tf(name):
    raise NotImplementedError(f"{name} inspects a TensorFlow graph and needs the TensorFlow runtime, which is not part of zoo-b200")


def find_tensors(sources, predicate):
    _needs_tf("find_tensors")


def find_placeholders(grads):
    _needs_tf("find_placeholders")


def to_bigdl_optim_method(koptim_method):
    """dict of optimizers -> dict; zoo OptimMethod / torch optimizer / optimizer name -> the engine's `OptConfig`
    (the reference maps tf.keras optimizers to BigDL OptimMethods with the same hyper-parameters)."""
    from ..keras.optimizers import to_opt_config

    if isinstance(koptim_method, dict):
        return {k: to_bigdl_optim_method(v) for k, v in koptim_method.items()}
    inner = getattr(koptim_method, "optimizer", None)
    if inner is not None and not isinstance(koptim_method, torch.optim.Optimizer):
        koptim_method = inner
    return to_opt_config(koptim_method)
