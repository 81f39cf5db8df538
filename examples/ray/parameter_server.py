"""Synchronous parameter server on RayContext actors (reference examples: pyzoo/zoo/examples/ray/parameter_server/
sync_parameter_server.py -- one ParameterServer actor, N DataWorker actors computing gradients of a small model)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.ray import RayContext


class ParameterServer:
    def __init__(self, dim, lr):
        self.w, self.lr = np.zeros(dim, np.float32), lr

    def apply_gradients(self, *grads):
        self.w -= self.lr * np.mean(grads, axis=0)
        return self.w

    def get_weights(self):
        return self.w


class DataWorker:
    def __init__(self, seed, dim):
        rng = np.random.RandomState(seed)
        self.x = rng.randn(512, dim).astype(np.float32)
        true_w = np.arange(1, dim + 1, dtype=np.float32)
        self.y = self.x @ true_w + 0.01 * rng.randn(512).astype(np.float32)

    def compute_gradients(self, w):
        err = self.x @ w - self.y
        return (self.x.T @ err / len(err)).astype(np.float32)


if __name__ == "__main__":
    dim, workers = 4, 2
    ctx = RayContext(sc=None, object_store_memory="1g").init()
    ps = ctx.actor(ParameterServer, dim, 0.3)
    ws = [ctx.actor(DataWorker, seed, dim) for seed in range(workers)]
    w = ps.get_weights.remote().get()
    for it in range(int(os.environ.get("ITERS", 60))):
        grads = [wk.compute_gradients.remote(w) for wk in ws]
        w = ps.apply_gradients.remote(*RayContext.get_objects(grads)).get()
    print("weights after training:", np.round(w, 2).tolist(), "(target [1, 2, 3, 4])")
    ctx.stop()
