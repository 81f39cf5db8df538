"""Data-parallel gradient exchange strategies.

The reference has one strategy: BigDL `AllReduceParameter` -- gradient slices are put into the
Spark BlockManager, slice *n* is summed and optimizer-stepped by task *n*, and every task fetches
all updated weight slices (Topology.scala:1128-1206, docs/docs/wp-bigdl.md:140-158).  The same
owner-computes schedule is kept, executed by ONE kernel over NVLink peer memory:

``P2PFusedSync``   reduce-scatter (peer loads) -> optimizer on the owned shard -> all-gather
                   (peer stores), `csrc/comm/rs_opt_ag.cu`; no NCCL call on the hot path.
``AllReduceSync``  `torch.distributed.all_reduce` (NCCL on GPUs, gloo on CPU) followed by the local
                   fused optimizer pass -- the *baseline* path, and the CPU plumbing path.
``LocalSync``      world_size == 1.
"""
from __future__ import annotations

import logging

import torch

log = logging.getLogger("zoo.dp")


class LocalSync:
    name = "local"

    def __init__(self, ctx):
        self.ctx = ctx
        self.fopt = None

    def allocator(self):
        return None

    def pad_to(self):
        return 1

    def attach(self, fopt):
        self.fopt = fopt

    def broadcast_parameters(self):
        pass

    def begin_step(self):
        """Start of a training step (before forward); see FusedOptimizer.begin_step."""
        self.fopt.begin_step()

    def step(self):
        self.fopt.step()

    def graph_safe(self):
        return True

    def close(self):
        pass


class AllReduceSync(LocalSync):
    """Library collective + local optimizer (baseline / CPU path)."""
    name = "allreduce"

    def attach(self, fopt):
        self.fopt = fopt
        fopt.set_grad_scale(1.0 / self.ctx.world_size)

    def broadcast_parameters(self):
        import torch.distributed as dist

        dist.broadcast(self.fopt.arena.flat_p, src=0)
        if self.fopt.arena.flat_p16 is not None:
            self.fopt.arena.flat_p16.copy_(self.fopt.arena.flat_p)
        # buffers (e.g. BatchNorm statistics) start identical too
        for b in self.fopt_buffers():
            dist.broadcast(b, src=0)

    def fopt_buffers(self):
        return []

    def graph_safe(self):
        return False

    def step(self):
        import torch.distributed as dist

        dist.all_reduce(self.fopt.arena.flat_g)
        self.fopt.step()


def make_grad_sync(ctx, mode="auto"):
    if ctx.world_size <= 1:
        return LocalSync(ctx)
    if mode == "auto":
        mode = ctx.conf.get("zoo.comm.backend", "auto")
    if mode == "auto":
        mode = "p2p" if ctx.is_gpu else "gloo"
    if mode in ("nccl", "gloo", "allreduce"):
        return AllReduceSync(ctx)
    if mode == "p2p":
        # No silent fallback: a job that asked for (or defaulted to) the fused peer-memory path and cannot have it must not
        # quietly run on the library all-reduce and still report itself as the product path.  Select the baseline explicitly
        # with ZOO_COMM_BACKEND=nccl / conf zoo.comm.backend=nccl.
        from .p2p import P2PFusedSync

        try:
            return P2PFusedSync(ctx)
        except Exception as e:
            raise RuntimeError(f"the fused P2P gradient path could not be set up ({e!r}); set ZOO_COMM_BACKEND=nccl to run "
                               "on the NCCL all-reduce baseline instead") from e
    raise ValueError(f"unknown gradient-exchange backend {mode!r} (auto | p2p | nccl | gloo)")
