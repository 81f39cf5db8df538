"""MXNetTrainer (reference: pyzoo/zoo/ray/mxnet/mxnet_trainer.py:26-139, mxnet_runner.py:28-195).

The reference runs MXNet workers and parameter servers as Ray actors with `kvstore=dist_sync`.  MXNet is
not part of this image; the *shape* of the API -- creator functions shipped to workers, synchronous
data-parallel training, per-worker stats back -- lives on as
`zoo.orca.learn.pytorch.Estimator.from_torch(model_creator=..., workers_per_node=N)`, whose gradient
exchange (sync parameter server == all-reduce) is the fused kernel K1.  This class keeps the constructor
and forwards torch creators to that path; MXNet symbols/gluon blocks are rejected with a clear message.
"""
from __future__ import annotations


class MXNetTrainer:
    def __init__(self, config, data_creator, model_creator, loss_creator=None, metrics_creator=None, num_workers=1,
                 num_servers=None, runner_cores=None, optimizer_creator=None):
        import torch

        self.config = dict(config or {})
        probe = model_creator(self.config)
        if not isinstance(probe, torch.nn.Module):
            raise NotImplementedError("MXNet models cannot run here (no mxnet in this image); pass torch creator functions "
                                      "or use zoo.orca.learn.pytorch.Estimator.from_torch")
        from ...orca.learn.pytorch import Estimator

        opt = optimizer_creator or (lambda m, c: torch.optim.SGD(m.parameters(), lr=c.get("lr", c.get("learning_rate", 0.01))))
        self.data_creator = data_creator
        self.num_workers = num_workers
        self.est = Estimator.from_torch(model_creator=model_creator, optimizer_creator=opt, loss_creator=loss_creator,
                                        metrics=metrics_creator(self.config) if metrics_creator else None, config=self.config,
                                        workers_per_node=num_workers)

    def train(self, nb_epoch=1):
        bs = int(self.config.get("batch_size", 32)) * max(1, self.num_workers)
        r = self.est.fit(self.data_creator, epochs=nb_epoch, batch_size=bs)
        return r if isinstance(r, list) else [r]

    def shutdown(self):
        self.est.shutdown()
