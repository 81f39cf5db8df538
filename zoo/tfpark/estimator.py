"""`zoo.tfpark.estimator.TFEstimator` (reference: pyzoo/zoo/tfpark/estimator.py:30-332).

The reference wraps `tf.estimator.Estimator`: `model_fn(features, labels, mode, params)` builds a graph and returns an
EstimatorSpec; `train / evaluate / predict` rebuild the graph for their mode, restore the variables from `model_dir`, and run it
on a TFDataset returned by `input_fn(mode)`; the train op must come from a `ZooOptimizer` (estimator.py:75-78).

Here the graph is a torch module, so `model_fn(features, labels, mode, params)` is called once per `train / evaluate / predict`
call -- `features` / `labels` are the first batch, which lets the function size its layers -- and returns a `TFEstimatorSpec`:

    def model_fn(features, labels, mode, params):
        net = nn.Sequential(nn.Linear(features.shape[1], 32), nn.ReLU(), nn.Linear(32, params["classes"]))
        return TFEstimatorSpec(mode, model=net, loss=nn.CrossEntropyLoss(),
                               train_op=ZooOptimizer(Adam(1e-3)), predictions=lambda out: out.argmax(-1))

As with tf.estimator, state lives in `model_dir`: every call restores the newest weights written there and `train` writes them
back, so the three calls can come from different processes.
"""
from __future__ import annotations

import os
import tempfile

import numpy as np
import torch
import torch.nn as nn

from ..common.triggers import MaxEpoch, MaxIteration
from ..pipeline.api.keras import metrics as _metrics
from .tf_dataset import TFDataset


class ModeKeys:
    TRAIN, EVAL, PREDICT = "train", "eval", "infer"


class TFEstimatorSpec:
    """What `model_fn` returns.  model: torch module (features -> outputs); loss: callable (outputs, labels) -> scalar;
    train_op: a ZooOptimizer / zoo optimizer / torch optimizer factory; predictions: optional callable mapping the model outputs
    to what `predict` yields (a tensor or a dict of tensors)."""

    def __init__(self, mode, predictions=None, loss=None, train_op=None, model=None, eval_metric_ops=None):
        if model is None or not isinstance(model, nn.Module):
            raise ValueError("TFEstimatorSpec needs model=<torch module>")
        self.mode, self.predictions, self.loss, self.train_op, self.model, self.eval_metric_ops = mode, predictions, loss, train_op, model, eval_metric_ops


class TFEstimator:
    def __init__(self, model_fn, model_dir=None, config=None, params=None, warm_start_from=None):
        if not callable(model_fn):
            raise TypeError("TFEstimator(estimator) wraps a tf.estimator.Estimator, which needs TensorFlow; use "
                            "TFEstimator.from_model_fn(model_fn, ...) with a model_fn returning a TFEstimatorSpec")
        self.model_fn, self.config, self.params = model_fn, config, dict(params or {})
        self.model_dir = model_dir or tempfile.mkdtemp(prefix="zoo_tfestimator_")
        self.warm_start_from = warm_start_from
        self.global_step = 0

    @classmethod
    def from_model_fn(cls, model_fn, model_dir=None, config=None, params=None, warm_start_from=None):
        return cls(model_fn, model_dir, config, params, warm_start_from)

    # ---- state in model_dir ------------------------------------------------------------------------------------------------
    def _ckpt(self):
        return os.path.join(self.model_dir, "model.ckpt.pt")

    def latest_checkpoint(self):
        return self._ckpt() if os.path.exists(self._ckpt()) else None

    def _restore(self, net, checkpoint_path=None):
        path = checkpoint_path or self.latest_checkpoint() or self.warm_start_from
        if path is not None and os.path.exists(path):
            blob = torch.load(path, map_location="cpu", weights_only=False)
            net.load_state_dict(blob["model"] if isinstance(blob, dict) and "model" in blob else blob)
            if isinstance(blob, dict) and path != self.warm_start_from:
                self.global_step = int(blob.get("global_step", self.global_step))

    def _save(self, net):
        os.makedirs(self.model_dir, exist_ok=True)
        tmp = self._ckpt() + ".tmp"
        torch.save({"model": net.state_dict(), "global_step": self.global_step}, tmp)
        os.replace(tmp, self._ckpt())

    # ---- model_fn plumbing -------------------------------------------------------------------------------------------------
    @staticmethod
    def _first_batch(ds):
        fs = ds.feature_set
        one = lambda t: None if t is None else ((t[0] if len(t) == 1 else list(t)) if isinstance(t, (tuple, list)) else t)
        for mb in fs.batches(2, train=False):
            return one(mb.inputs), one(mb.targets)
        raise ValueError("input_fn returned an empty dataset")

    def _call_model_fn(self, ds, mode):
        import inspect

        x, y = self._first_batch(ds)
        if mode == ModeKeys.PREDICT:
            y = None
        names = inspect.signature(self.model_fn).parameters
        kw = {}
        if "mode" in names:
            kw["mode"] = mode
        if "params" in names:
            kw["params"] = self.params
        if "config" in names:
            kw["config"] = self.config
        spec = self.model_fn(x, y, **kw)
        if not isinstance(spec, TFEstimatorSpec):
            raise TypeError("model_fn must return a zoo.tfpark.estimator.TFEstimatorSpec")
        return spec

    @staticmethod
    def _dataset(input_fn, mode):
        import inspect

        ds = input_fn(mode) if len(inspect.signature(input_fn).parameters) >= 1 else input_fn()
        if not isinstance(ds, TFDataset):
            raise TypeError("input_fn must return a TFDataset")
        return ds

    # ---- the three calls ---------------------------------------------------------------------------------------------------
    def train(self, input_fn, steps=None):
        from ..pipeline.estimator import Estimator
        from .unsupported import ZooOptimizer

        ds = self._dataset(input_fn, ModeKeys.TRAIN)
        if ds.batch_size <= 0:
            raise ValueError("You should set batch_size instead of batch_per_thread for training")
        spec = self._call_model_fn(ds, ModeKeys.TRAIN)
        if spec.loss is None or spec.train_op is None:
            raise ValueError("in TRAIN mode the spec needs loss= and train_op= (a ZooOptimizer)")
        self._restore(spec.model)
        opt = spec.train_op.optimizer if isinstance(spec.train_op, ZooOptimizer) else spec.train_op
        est = Estimator(spec.model, opt, None)
        end = MaxIteration(steps) if steps is not None else MaxEpoch(1)
        est.train(ds.feature_set, spec.loss, end, None, None, None, ds.batch_size)
        self.global_step += est.engine.state["neval"] - 1
        self.last_loss = est.engine.state.get("Loss")
        self._save(spec.model)
        est.close()
        return self

    def evaluate(self, input_fn, eval_methods, steps=None, checkpoint_path=None):
        from ..pipeline.estimator import Estimator

        ds = self._dataset(input_fn, ModeKeys.EVAL)
        spec = self._call_model_fn(ds, ModeKeys.EVAL)
        self._restore(spec.model, checkpoint_path)
        bs = ds.batch_per_thread if ds.batch_per_thread > 0 else ds.batch_size
        est = Estimator(spec.model, None, None)
        methods = [_metrics.to_metric(m, spec.loss) for m in eval_methods]
        res = est.evaluate(ds.feature_set, methods, bs, spec.loss)
        est.close()
        return {(m if isinstance(m, str) else r.name): r.value for m, r in zip(eval_methods, res)}

    def predict(self, input_fn, predict_keys=None, checkpoint_path=None):
        ds = self._dataset(input_fn, ModeKeys.PREDICT)
        spec = self._call_model_fn(ds, ModeKeys.PREDICT)
        self._restore(spec.model, checkpoint_path)
        dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        net = spec.model.to(dev).eval()
        bs = ds.batch_per_thread if ds.batch_per_thread > 0 else ds.batch_size
        outs = []
        with torch.no_grad():
            for mb in ds.feature_set.batches(max(bs, 1), train=False):
                xs = mb.inputs if isinstance(mb.inputs, (tuple, list)) else (mb.inputs,)
                out = net(*[t.to(dev) for t in xs])
                p = spec.predictions(out) if callable(spec.predictions) else out
                if isinstance(p, dict):
                    keys = predict_keys or list(p.keys())
                    n = len(next(iter(p.values())))
                    outs.extend({k: p[k][i].cpu().numpy() for k in keys} for i in range(n))
                else:
                    outs.extend(np.asarray(p.float().cpu() if p.is_floating_point() else p.cpu()))
        return outs
