"""MXNetRunner (reference: pyzoo/zoo/ray/mxnet/mxnet_runner.py:28-195): the Ray actor that is one MXNet worker or parameter
server (`setup_distributed(env, config, data_creator, model_creator, ...)`, `train(nb_epoch)`, `shutdown`, `get_node_ip`,
`find_free_port`).  Synchronous parameter servers are the all-reduce here, so there is no server role: a runner is one
data-parallel torch worker of the local pool, driven by `MXNetTrainer`."""
from __future__ import annotations

import socket

from .utils import find_free_port


class MXNetRunner:
    def setup_distributed(self, env, config, data_creator, model_creator, loss_creator=None, metrics_creator=None):
        import torch

        self.env, self.config = dict(env or {}), dict(config or {})
        self.is_worker = self.env.get("DMLC_ROLE", "worker") == "worker"
        if not self.is_worker:
            return self          # parameter-server / scheduler roles have nothing to do: gradients are all-reduced
        model = model_creator(self.config)
        if not isinstance(model, torch.nn.Module):
            raise NotImplementedError("MXNet models cannot run here (no mxnet in this image); pass torch creator functions")
        self.model = model
        self.loss = loss_creator(self.config) if loss_creator else torch.nn.CrossEntropyLoss()
        self.metrics = metrics_creator(self.config) if metrics_creator else None
        self.data = data_creator(self.config, self.env)
        lr = self.config.get("optimizer_params", {}).get("learning_rate", 0.01)
        self.optimizer = torch.optim.SGD(self.model.parameters(), lr=lr)
        self.epoch = 0
        return self

    def train(self, nb_epoch=1):
        import torch

        stats = {}
        if not getattr(self, "is_worker", False):
            return stats
        train_data = self.data[0] if isinstance(self.data, tuple) else self.data      # data_creator may return (train, validation)
        for _ in range(nb_epoch):
            tot, n = 0.0, 0
            for x, y in train_data:
                loss = self.loss(self.model(torch.as_tensor(x)), torch.as_tensor(y))
                self.optimizer.zero_grad()
                loss.backward()
                self._exchange_gradients()
                self.optimizer.step()
                tot, n = tot + float(loss.detach()), n + 1
            self.epoch += 1
            stats = {"epoch": self.epoch, "train_loss": tot / max(n, 1)}
        return stats

    def _exchange_gradients(self):
        """`kvstore=dist_sync` semantics: every worker applies the mean gradient.  Runners started inside a process group (the
        pool sets RANK / WORLD_SIZE) average through it; a lone runner has nothing to exchange."""
        import torch.distributed as dist

        if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
            return
        W = dist.get_world_size()
        for p in self.model.parameters():
            if p.grad is not None:
                dist.all_reduce(p.grad)
                p.grad.div_(W)

    def shutdown(self):
        for k in ("model", "data", "optimizer"):
            self.__dict__.pop(k, None)

    def get_node_ip(self):
        try:
            return socket.gethostbyname(socket.gethostname())
        except OSError:
            return "127.0.0.1"

    def find_free_port(self):
        return find_free_port()
