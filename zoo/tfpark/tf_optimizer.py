"""`zoo.tfpark.tf_optimizer` (reference: pyzoo/zoo/tfpark/tf_optimizer.py:40-709).

The reference's TFOptimizer exports a TensorFlow training graph (loss, gradients, variables) into a `TFTrainingHelper` layer and
lets the BigDL distributed optimizer drive it.  The training graph here is a torch module, and the driver is this framework's
engine (flat fp32 master arena, fused optimizer kernels, P2P gradient exchange), so the same three objects exist with torch content:

  * `TFTrainingHelper(model, loss)`  -- a module whose forward returns the LOSS (and extra metric outputs), i.e. "the graph with
    the loss inside" (tf_optimizer.py:68-96); weights are readable / checkpointable through it.
  * `TFModel`                        -- (training helper, criterion, validation methods) (tf_optimizer.py:106-340).
  * `TFOptimizer`                    -- `from_loss`, `from_keras`, `optimize(end_trigger, checkpoint_trigger)`, gradient
    clipping, summaries, `load_checkpoint` (tf_optimizer.py:342-709).  `from_train_op` takes a TensorFlow op and stays TF-only.

The metric adapters keep their meaning: `StatelessMetric(name, idx)` averages the idx-th extra output of the model,
`BigDLMetric` wraps a zoo validation method.
"""
from __future__ import annotations

import os

import torch
import torch.nn as nn

from ..common.triggers import EveryEpoch, MaxEpoch
from ..pipeline.api.keras import metrics as _metrics


class TFTrainingHelper(nn.Module):
    """model + loss as one module: `forward(*inputs_and_labels)` -> loss (scalar).  `n_inputs` says how many leading tensors are
    model inputs; the rest are labels."""

    def __init__(self, model, loss, n_inputs=1, metrics=None):
        super().__init__()
        self.model, self.loss_fn, self.n_inputs = model, loss, int(n_inputs)
        self.metric_fns = dict(metrics or {})

    def forward(self, *tensors):
        xs, ys = tensors[:self.n_inputs], tensors[self.n_inputs:]
        out = self.model(*xs)
        y = ys[0] if len(ys) == 1 else (list(ys) if ys else None)
        loss = self.loss_fn(out, y) if y is not None else self.loss_fn(out)
        if not self.metric_fns:
            return loss
        return [loss] + [f(out, y) for f in self.metric_fns.values()]

    def get_weights_to_python(self):
        return [p.detach().cpu().numpy() for p in self.model.parameters()]

    def save_checkpoint(self, path=None):
        if path is not None:
            torch.save(self.model.state_dict(), path)
        return self

    def load_checkpoint(self, path):
        self.model.load_state_dict(torch.load(path, map_location="cpu"))
        return self


class TFModel:
    """(training_helper_layer, criterion, val_methods)."""

    def __init__(self, training_helper_layer, criterion=None, val_methods=None):
        self.training_helper_layer = training_helper_layer
        self.criterion = criterion if criterion is not None else IdentityCriterion()
        self.val_methods = val_methods

    @staticmethod
    def create(loss, model=None, inputs=1, metrics=None, val_methods=None, **kw):
        """`loss`: callable (outputs, labels) -> scalar; `model`: the torch module; `inputs`: number of model inputs."""
        if model is None or not callable(loss):
            raise NotImplementedError("TFModel.create(loss_tensor, sess, ...) exports a TensorFlow graph; here pass a loss callable and "
                                      "model=<torch module>")
        return TFModel(TFTrainingHelper(model, loss, inputs if isinstance(inputs, int) else len(inputs), metrics), IdentityCriterion(), val_methods)

    create_for_unfreeze = create


class IdentityCriterion:
    """The loss is computed inside the model graph; the criterion returns its first output unchanged."""

    def __call__(self, output, target=None):
        out = output[0] if isinstance(output, (list, tuple)) else output
        return out.mean() if hasattr(out, "mean") else out


class TFValidationMethod:
    """(val_method, name, output_indices, label_indices): picks outputs / labels by index and applies `val_method`."""

    def __init__(self, val_method, name, output_indices, label_indices):
        self.val_method, self.name, self.output_indices, self.label_indices = val_method, name, list(output_indices), list(label_indices)

    def __call__(self, outputs, labels):
        pick = lambda xs, idx: [xs[i] for i in idx] if isinstance(xs, (list, tuple)) else xs
        o, l = pick(outputs, self.output_indices), pick(labels, self.label_indices)
        return self.val_method(o[0] if isinstance(o, list) and len(o) == 1 else o, l[0] if isinstance(l, list) and len(l) == 1 else l)


class StatelessMetric:
    """Mean over batches of the `idx`-th metric tensor the model produces."""

    def __init__(self, metric_name, idx):
        self.name, self.idx, self.total, self.count = metric_name, idx, 0.0, 0

    def update(self, outputs, batch_size=1):
        v = outputs[self.idx] if isinstance(outputs, (list, tuple)) else outputs
        self.total += float(v.mean() if hasattr(v, "mean") else v) * batch_size
        self.count += batch_size
        return self

    def result(self):
        return self.total / max(self.count, 1), self.count


class BigDLMetric:
    def __init__(self, val_method, outputs=None, labels=None):
        self.val_method = _metrics.to_metric(val_method) if hasattr(_metrics, "to_metric") else val_method
        self.outputs, self.labels = outputs, labels


class TFOptimizer:
    """Distributed training of a loss over a TFDataset (tf_optimizer.py:342-709).

        opt = TFOptimizer.from_loss(loss_fn, Adam(1e-3), model=net, dataset=TFDataset.from_ndarrays((x, y), batch_size=256))
        opt.optimize(end_trigger=MaxEpoch(5))

    `from_keras(model, dataset)` takes a compiled zoo Keras-API model / `tfpark.KerasModel`.  Everything runs on the engine that
    `Estimator.from_torch` uses, so data parallelism, checkpoints, summaries and clipping behave the same."""

    def __init__(self, tf_model, optim_method, sess=None, dataset=None, val_split=0.0, clip_norm=None, clip_value=None, metrics=None,
                 tensor_with_value=None, session_config=None, model_dir=None, updates=None):
        from ..pipeline.estimator import Estimator
        from .unsupported import ZooOptimizer

        if not isinstance(tf_model, TFModel):
            raise TypeError("TFOptimizer takes a TFModel; use TFOptimizer.from_loss / from_keras")
        if dataset is None:
            raise ValueError("dataset (a TFDataset with batch_size set) is required")
        if dataset.batch_size <= 0:
            raise ValueError("You should set batch_size instead of batch_per_thread for training")
        if isinstance(optim_method, ZooOptimizer):
            optim_method = optim_method.optimizer
        self.tf_model, self.dataset, self.optim_method, self.model_dir = tf_model, dataset, optim_method, model_dir
        self.val_split = val_split
        helper = tf_model.training_helper_layer
        self._net, self._loss = helper.model, helper.loss_fn
        self.estimator = Estimator(self._net, optim_method, model_dir)
        if clip_norm is not None:
            self.estimator.set_l2_norm_gradient_clipping(clip_norm)
        if clip_value is not None:
            lo, hi = clip_value if isinstance(clip_value, (tuple, list)) else (-abs(clip_value), abs(clip_value))
            self.estimator.set_constant_gradient_clipping(lo, hi)
        self.metrics = metrics

    # ---- constructors ----------------------------------------------------------------------------------------------
    @classmethod
    def from_loss(cls, loss, optim_method, session=None, val_outputs=None, val_labels=None, val_method=None, val_split=0.0,
                  clip_norm=None, clip_value=None, metrics=None, tensor_with_value=None, session_config=None, model_dir=None,
                  updates=None, model=None, dataset=None, inputs=1):
        tfm = TFModel.create(loss, model=model, inputs=inputs, val_methods=val_method)
        return cls(tfm, optim_method, session, dataset, val_split, clip_norm, clip_value, metrics, tensor_with_value, session_config,
                   model_dir, updates)

    @classmethod
    def from_keras(cls, keras_model, dataset, optim_method=None, val_split=0.0, session_config=None, model_dir=None, **kw):
        from .model import KerasModel

        net = keras_model.model if isinstance(keras_model, KerasModel) else keras_model
        comp = getattr(net, "_compiled", None)
        if comp is None:
            raise ValueError("from_keras needs a compiled model: model.compile(optimizer, loss, metrics)")
        tfm = TFModel(TFTrainingHelper(net, comp["loss"]), IdentityCriterion(), comp["metrics"] or None)
        return cls(tfm, optim_method if optim_method is not None else comp["optimizer"], None, dataset, val_split, model_dir=model_dir)

    @classmethod
    def from_train_op(cls, train_op, loss, **kw):
        raise NotImplementedError("TFOptimizer.from_train_op takes a TensorFlow train op; there is no TensorFlow runtime in zoo-b200. "
                                  "Use TFOptimizer.from_loss(loss_fn, optim_method, model=..., dataset=...)")

    @staticmethod
    def export_training_model(export_dir, loss, sess=None, inputs=1, labels=None, predictions=None, metrics=None, tensor_with_value=None,
                              updates=None, model=None):
        """Write the training model (weights + structure) under `export_dir` (reference: tf_optimizer.py:568-576)."""
        if model is None:
            raise NotImplementedError("export_training_model(loss_tensor, sess, ...) exports a TensorFlow graph; pass model=<torch module>")
        os.makedirs(export_dir, exist_ok=True)
        torch.save(TFTrainingHelper(model, loss, inputs if isinstance(inputs, int) else len(inputs)), os.path.join(export_dir, "training_model.pt"))
        return export_dir

    # ---- configuration -----------------------------------------------------------------------------------------------
    def set_constant_gradient_clipping(self, min_value, max_value):
        self.estimator.set_constant_gradient_clipping(min_value, max_value)

    def set_gradient_clipping_by_l2_norm(self, clip_norm):
        self.estimator.set_l2_norm_gradient_clipping(clip_norm)

    def set_train_summary(self, summary):
        self.estimator.engine.train_summary = summary

    def set_val_summary(self, summary):
        self.estimator.engine.val_summary = summary

    def load_checkpoint(self, path, version=None):
        """Resume from `path/model.<version>` written by `optimize(checkpoint_trigger=...)`."""
        eng = self.estimator.engine
        if version is not None:
            f = os.path.join(path, f"model.{version}")
        else:
            files = eng._model_files(path)
            if not files:
                raise FileNotFoundError(f"no model.<n> checkpoint under {path}")
            f = max(files, key=lambda q: int(q.rsplit(".", 1)[-1]))
        eng.load_checkpoint(f)
        return self

    # ---- run -------------------------------------------------------------------------------------------------------------
    def optimize(self, end_trigger=None, checkpoint_trigger=None):
        end_trigger = end_trigger if end_trigger is not None else MaxEpoch(1)
        ds = self.dataset
        val = ds.val_set
        methods = self.tf_model.val_methods or self.metrics
        if isinstance(methods, dict):
            methods = list(methods.values())
        if val is not None and not methods:
            methods = ["loss"]
        if checkpoint_trigger is None and (val is not None or self.model_dir):
            checkpoint_trigger = EveryEpoch()
        self.estimator.train(ds.feature_set, self._loss, end_trigger, checkpoint_trigger, val, methods if val is not None else None, ds.batch_size)
        return self
