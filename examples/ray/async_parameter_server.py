"""Asynchronous parameter server on RayContext actors (reference example: pyzoo/zoo/examples/ray/parameter_server/
async_parameter_server.py -- workers pull the current weights, compute a gradient and push it; the server applies every
gradient as it arrives, without waiting for the other workers)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.ray import RayContext


class ParameterServer:
    def __init__(self, dim, lr):
        self.w, self.lr, self.updates = np.zeros(dim, np.float32), lr, 0

    def push(self, grad):
        self.w -= self.lr * grad
        self.updates += 1
        return self.updates

    def pull(self):
        return self.w

    def stats(self):
        return self.updates


class Worker:
    def __init__(self, seed, dim):
        rng = np.random.RandomState(seed)
        self.x = rng.randn(512, dim).astype(np.float32)
        self.y = self.x @ np.arange(1, dim + 1, dtype=np.float32)

    def gradient(self, w, it):
        rows = np.arange(it * 64, it * 64 + 64) % len(self.x)           # a mini-batch per call
        err = self.x[rows] @ w - self.y[rows]
        return (self.x[rows].T @ err / len(rows)).astype(np.float32)


if __name__ == "__main__":
    dim, n_workers, iters = 4, 3, int(os.environ.get("ITERS", 60))
    ctx = RayContext(sc=None, object_store_memory="1g").init()
    ps = ctx.actor(ParameterServer, dim, 0.1)
    workers = [ctx.actor(Worker, seed, dim) for seed in range(n_workers)]
    # every worker runs its own pull -> gradient -> push chain; chains of different workers interleave freely on the server
    pending = {i: w.gradient.remote(ps.pull.remote().get(), 0) for i, w in enumerate(workers)}
    done = {i: 0 for i in pending}
    while any(v < iters for v in done.values()):
        for i, ref in list(pending.items()):
            if done[i] >= iters:
                continue
            ps.push.remote(ref.get()).get()
            done[i] += 1
            if done[i] < iters:
                pending[i] = workers[i].gradient.remote(ps.pull.remote().get(), done[i])
    print("updates applied:", ps.stats.remote().get(), "weights:", np.round(ps.pull.remote().get(), 2).tolist(), "(target [1, 2, 3, 4])")
    ctx.stop()
