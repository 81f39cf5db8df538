"""`zoo.orca.learn.pytorch.Estimator.from_torch` -- creator-function estimator that can also spawn its
own workers.

Inside a launched job (torchrun, or one process) this is `zoo.pipeline.estimator.Estimator.from_torch`.
With `workers_per_node=N > 1` from a plain Python process it plays the role of the reference's
`MXNetTrainer` (pyzoo/zoo/ray/mxnet/mxnet_trainer.py:63-139): the creators are shipped to N spawned
workers (one per GPU; gloo on CPU), every worker builds model/optimizer/data from them, trains
data-parallel with the fused gradient exchange, and the driver gets rank 0's stats and weights back.
"""
from __future__ import annotations

import os

import torch

from ....pipeline.estimator.estimator import Estimator as _Base
from ....pipeline.estimator.estimator import TorchEstimator


class _SpawningEstimator:
    def __init__(self, model_creator, optimizer_creator, loss_creator, metrics, config, workers, model_dir):
        if model_creator is None or not callable(model_creator):
            raise ValueError("workers_per_node > 1 needs creator functions (model_creator, optimizer_creator), like MXNetTrainer")
        self.c = dict(model_creator=model_creator, optimizer_creator=optimizer_creator, loss_creator=loss_creator,
                      metrics=metrics, config=dict(config or {}), model_dir=model_dir)
        self.workers = int(workers)
        self.state = None                     # rank-0 model state_dict after the last fit
        self.model = model_creator(self.c["config"])

    def _job(self, what, data_creator, kw):
        c, state = self.c, self.state

        def run():
            from zoo.pipeline.estimator.estimator import TorchEstimator as TE

            est = TE(model_creator=c["model_creator"], optimizer_creator=c["optimizer_creator"], loss_creator=c["loss_creator"],
                     metrics=c["metrics"], config=c["config"], model_dir=c["model_dir"])
            if state is not None:
                est.model.load_state_dict(state)
            data = data_creator(c["config"], kw.get("batch_size", 32)) if callable(data_creator) else data_creator
            if what == "fit":
                stats = est.fit(data, **kw)
                sd = {k: v.detach().cpu() for k, v in est.model.state_dict().items()}
                return stats, sd
            if what == "evaluate":
                return est.evaluate(data, **kw), None
            return est.predict(data, **kw), None

        from zoo.util.launcher import launch

        return launch(run, self.workers)

    def fit(self, data, epochs=1, batch_size=32, **kw):
        res = self._job("fit", data, dict(epochs=epochs, batch_size=batch_size, **kw))
        self.state = res[0][1]
        self.model.load_state_dict(self.state)
        return [r[0] for r in res]

    def evaluate(self, data, batch_size=32, **kw):
        return self._job("evaluate", data, dict(batch_size=batch_size, **kw))[0][0]

    def predict(self, data, batch_size=32, **kw):
        return self._job("predict", data, dict(batch_size=batch_size, **kw))[0][0]

    def get_model(self):
        return self.model

    def save(self, path):
        torch.save({"model": self.state or self.model.state_dict()}, path)
        return path

    def load(self, path):
        self.state = torch.load(path, map_location="cpu", weights_only=False)["model"]
        self.model.load_state_dict(self.state)
        return self

    def shutdown(self):
        pass


class Estimator(_Base):
    @staticmethod
    def from_torch(*, model=None, optimizer=None, loss=None, metrics=None, model_dir=None, config=None, model_creator=None,
                   optimizer_creator=None, loss_creator=None, backend="b200", workers_per_node=1, ctx=None, **_ignored):
        if callable(model) and not isinstance(model, torch.nn.Module):
            model_creator, model = model, None
        if callable(optimizer) and not isinstance(optimizer, (torch.optim.Optimizer, type)):
            optimizer_creator, optimizer = optimizer, None
        in_job = int(os.environ.get("WORLD_SIZE", "1")) > 1
        if workers_per_node > 1 and not in_job:
            lc = loss_creator if loss_creator is not None else ((lambda cfg, _l=loss: _l) if loss is not None else None)
            return _SpawningEstimator(model_creator, optimizer_creator, lc, metrics, config, workers_per_node, model_dir)
        return TorchEstimator(model=model, optimizer=optimizer, loss=loss, metrics=metrics, model_dir=model_dir, config=config,
                              model_creator=model_creator, optimizer_creator=optimizer_creator, loss_creator=loss_creator,
                              backend=backend, ctx=ctx)
