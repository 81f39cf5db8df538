"""Estimator: train / evaluate a model on FeatureSets.

Reference API (pyzoo/zoo/pipeline/estimator/estimator.py:21-165, Estimator.scala:33-264):
``Estimator(model, optim_methods=None, model_dir=None)`` with ``train``, ``train_minibatch``,
``train_imagefeature``, ``evaluate``, ``evaluate_imagefeature``, ``evaluate_minibatch`` and the
three gradient-clipping setters.  `optim_methods` is one optimizer or a dict
``{sub_module_name: optimizer}`` (wide part FTRL + deep part Adagrad, CensusWideAndDeep.scala:122-132).

North-star additions: ``Estimator.from_torch(model|model_creator,
optimizer|optimizer_creator, loss, metrics, model_dir, backend)`` returning an estimator with
``fit / evaluate / predict / save / load / shutdown`` -- the creator-function shape of the
reference's `MXNetTrainer` (pyzoo/zoo/ray/mxnet/mxnet_trainer.py:63-132).
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.nn as nn

from ...common.nncontext import get_context
from ...common.triggers import EveryEpoch, MaxEpoch, MaxIteration, ZooTrigger
from ...feature.common import FeatureSet
from ..engine import TrainingEngine


class Estimator:
    def __init__(self, model, optim_methods=None, model_dir=None, bigdl_type="float", ctx=None):
        self.model = model
        self.optim_methods = optim_methods
        self.model_dir = model_dir
        self.engine = TrainingEngine(model, optim_methods, model_dir, ctx or get_context())
        self._loss = None
        self._metrics = None

    # ---- gradient clipping (estimator.py:34-59) ---------------------------------------
    def clear_gradient_clipping(self):
        self.engine.clear_gradient_clipping()

    def set_constant_gradient_clipping(self, min, max):
        self.engine.set_constant_gradient_clipping(min, max)

    def set_l2_norm_gradient_clipping(self, clip_norm):
        self.engine.set_gradient_clipping_by_l2_norm(clip_norm)

    set_gradient_clipping_by_l2_norm = set_l2_norm_gradient_clipping

    def set_drop_module_property(self, drop_percentage, max_drop_percentage, batch_size=100, warmup_iteration=200, micro_batches=4):
        """Gradient dropping / straggler mitigation (Topology.scala:1096-1100; `zoo.pipeline.straggler`)."""
        self.engine.set_drop_module_property(drop_percentage, max_drop_percentage, batch_size, warmup_iteration, micro_batches)
        return self

    def set_tensorboard(self, log_dir, app_name):
        self.engine.set_tensorboard(log_dir, app_name)

    def get_train_summary(self, tag=None):
        s = self.engine.train_summary
        return [] if s is None else s.read_scalar(tag)

    def get_validation_summary(self, tag=None):
        s = self.engine.val_summary
        return [] if s is None else s.read_scalar(tag)

    # ---- reference entry points -------------------------------------------------------
    def train(self, train_set, criterion, end_trigger=None, checkpoint_trigger=None, validation_set=None,
              validation_method=None, batch_size=32):
        """Train on a FeatureSet of Samples; `batch_size` is the global batch (estimator.py:61-81)."""
        self.engine.train(train_set, criterion, end_trigger, checkpoint_trigger, validation_set, validation_method,
                          batch_size)
        return self

    def train_minibatch(self, train_set, criterion, end_trigger=None, checkpoint_trigger=None, validation_set=None,
                        validation_method=None):
        """`train_set` already yields MiniBatches (e.g. a wrapped DataLoader) (estimator.py:83-104)."""
        bs = getattr(getattr(train_set, "_loader", None), "batch_size", None) or 32
        self.engine.train(train_set, criterion, end_trigger, checkpoint_trigger, validation_set, validation_method,
                          bs * self.engine.ctx.world_size)
        return self

    def train_imagefeature(self, train_set, criterion, end_trigger=None, checkpoint_trigger=None, validation_set=None,
                           validation_method=None, batch_size=32):
        return self.train(train_set, criterion, end_trigger, checkpoint_trigger, validation_set, validation_method, batch_size)

    def evaluate(self, validation_set, validation_method, batch_size=32, criterion=None):
        res = self.engine.evaluate(validation_set, validation_method, batch_size, criterion)
        return res

    evaluate_imagefeature = evaluate

    def evaluate_minibatch(self, validation_set, validation_method):
        bs = getattr(getattr(validation_set, "_loader", None), "batch_size", None) or 32
        return self.engine.evaluate(validation_set, validation_method, bs)

    def close(self):
        self.engine.close()

    # ---- Estimator.from_torch (north-star API) ---------------------------------------
    @staticmethod
    def from_torch(*, model=None, optimizer=None, loss=None, metrics=None, model_dir=None, config=None,
                   model_creator=None, optimizer_creator=None, loss_creator=None, backend="b200", workers_per_node=1,
                   ctx=None):
        return TorchEstimator(model=model, optimizer=optimizer, loss=loss, metrics=metrics, model_dir=model_dir,
                              config=config, model_creator=model_creator, optimizer_creator=optimizer_creator,
                              loss_creator=loss_creator, backend=backend, ctx=ctx)


class TorchEstimator:
    """`Estimator.from_torch(...)`: fit / evaluate / predict / save / load / shutdown."""

    def __init__(self, model=None, optimizer=None, loss=None, metrics=None, model_dir=None, config=None,
                 model_creator=None, optimizer_creator=None, loss_creator=None, backend="b200", ctx=None):
        config = dict(config or {})
        if model is None:
            if model_creator is None:
                raise ValueError("pass `model` or `model_creator`")
            model = model_creator(config)
        if optimizer is None and optimizer_creator is not None:
            optimizer = optimizer_creator(model, config)
        if loss is None and loss_creator is not None:
            loss = loss_creator(config) if callable(loss_creator) and not isinstance(loss_creator, nn.Module) else loss_creator
        if isinstance(loss, type):
            loss = loss()
        self.model, self.loss, self.metrics, self.config = model, loss, list(metrics or []), config
        self.engine = TrainingEngine(model, optimizer, model_dir, ctx or get_context())
        self.model_dir = model_dir

    def fit(self, data, epochs=1, batch_size=32, validation_data=None, checkpoint_trigger=None, feature_cols=None,
            label_cols=None, max_steps=None, log_every=None, shard=True):
        """data: FeatureSet | XShards | (x, y) ndarrays | DataLoader | data_creator(config, batch_size)."""
        if callable(data) and not isinstance(data, (FeatureSet, torch.utils.data.DataLoader)):
            data = data(self.config, batch_size)
        fs = self._to_set(data, feature_cols, label_cols)
        val = self._to_set(validation_data, feature_cols, label_cols) if validation_data is not None else None
        end = MaxIteration(self.engine.state["neval"] - 1 + max_steps) if max_steps else MaxEpoch(self.engine.state["epoch"] + epochs - 1)
        trig = checkpoint_trigger or (EveryEpoch() if (val is not None or self.model_dir) else None)
        self.engine.train(fs, self.loss, end, trig, val, self.metrics if val is not None else None, batch_size,
                          shard=shard, log_every=log_every)
        st = self.engine.state
        return {"epoch": st["epoch"] - 1, "iterations": st["neval"] - 1, "loss": st["Loss"],
                "throughput": self.engine.last_throughput}

    def _to_set(self, data, feature_cols=None, label_cols=None, batch_size=32):
        if data is None or isinstance(data, FeatureSet):
            return data
        if callable(data) and not isinstance(data, torch.utils.data.DataLoader) and not hasattr(data, "collect"):
            data = data(self.config, batch_size)          # a data_creator(config, batch_size)
            if isinstance(data, FeatureSet):
                return data
        if hasattr(data, "collect") and hasattr(data, "num_partitions"):
            return FeatureSet.xshards(data, feature_cols, label_cols)
        return self.engine._make_set(data)

    def evaluate(self, data, batch_size=32, feature_cols=None, label_cols=None):
        res = self.engine.evaluate(self._to_set(data, feature_cols, label_cols, batch_size), self.metrics or ["loss"], batch_size, self.loss)
        return {r.name: r.value for r in res}

    def predict(self, data, batch_size=32, feature_cols=None):
        if hasattr(data, "collect") and hasattr(data, "num_partitions"):
            data = FeatureSet.xshards(data, feature_cols, None)
        elif callable(data) and not isinstance(data, (FeatureSet, torch.utils.data.DataLoader)):
            data = data(self.config, batch_size)
            if isinstance(data, (tuple, list)) and len(data) == 2:
                data = data[0]
        return self.engine.predict(data, batch_size)

    def get_model(self):
        return self.model

    def save(self, path):
        """Model + optimizer state in one file (readable by `load`)."""
        self.engine.prepare()
        os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
        torch.save({"model": {k: v.detach().cpu() for k, v in self.model.state_dict().items()},
                    "optim": self.engine.fopt.state_dict(), "sparse": self.engine._sparse_state(),
                    "state": dict(self.engine.state)}, path)
        return path

    def load(self, path):
        self.engine.prepare()
        blob = torch.load(path, map_location="cpu", weights_only=False)
        with torch.no_grad():
            own = self.model.state_dict()
            for k, v in blob["model"].items():
                if k in own:
                    own[k].copy_(v)
        self.engine.fopt.load_state_dict(blob["optim"])
        self.engine.state.update(blob.get("state", {}))
        if self.engine.fopt.arena.flat_p16 is not None:
            self.engine.fopt.arena.flat_p16.copy_(self.engine.fopt.arena.flat_p)
        self.engine._after_step_shadows()
        return self

    def load_orca_checkpoint(self, path, version=None):
        self.engine.set_checkpoint(path, self.engine.overwrite_checkpoint)
        p = None if version is None else os.path.join(path, f"model.{version}")
        self.engine.load_checkpoint(p)
        return self

    def set_tensorboard(self, log_dir, app_name):
        self.engine.set_tensorboard(log_dir, app_name)

    def get_train_summary(self, tag=None):
        s = self.engine.train_summary
        return [] if s is None else s.read_scalar(tag)

    def get_validation_summary(self, tag=None):
        s = self.engine.val_summary
        return [] if s is None else s.read_scalar(tag)

    def set_constant_gradient_clipping(self, min, max):
        self.engine.set_constant_gradient_clipping(min, max)

    def set_l2_norm_gradient_clipping(self, clip_norm):
        self.engine.set_gradient_clipping_by_l2_norm(clip_norm)

    def clear_gradient_clipping(self):
        self.engine.clear_gradient_clipping()

    def set_drop_module_property(self, drop_percentage, max_drop_percentage, batch_size=100, warmup_iteration=200, micro_batches=4):
        """Gradient dropping / straggler mitigation (Topology.scala:1096-1100; `zoo.pipeline.straggler`)."""
        self.engine.set_drop_module_property(drop_percentage, max_drop_percentage, batch_size, warmup_iteration, micro_batches)
        return self

    def shutdown(self):
        self.engine.close()
