"""Where does the end-to-end gap come from?  Wall-clock per step of variants of the host loop on the NCF bench config.
usage: python benchmarks/e2e_variants.py [scale]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import build_ncf
from zoo.common.nncontext import init_nncontext
from zoo.feature.common import FeatureSet, MiniBatch
from zoo.pipeline.engine import Prefetcher
from zoo.pipeline.estimator import Estimator


class A:
    config, scale, batch = "ncf", float(sys.argv[1]) if len(sys.argv) > 1 else 1.0, 0


ctx = init_nncontext()
cfg = build_ncf(A, ctx, torch)
opt = torch.optim.Adam([p for p in cfg["model"].parameters() if p.requires_grad] or [torch.zeros(1, requires_grad=True)], lr=1e-3)
est = Estimator.from_torch(model=cfg["model"], optimizer=opt, loss=cfg["loss"], ctx=ctx)
eng = est.engine
eng.prepare()
crit = cfg["loss"].to(ctx.device)
host = cfg["make_batches"](48, torch.device("cpu"))
X, Y = torch.cat([b[0] for b in host]), torch.cat([b[1] for b in host])
fs = FeatureSet(features=X, labels=Y, memory_type="DRAM", shuffle=False)
fs.cache()
print("pinned:", fs.features[0].is_pinned(), fs.labels[0].is_pinned())
B = cfg["batch"]
dev_batches = [MiniBatch(X[i * B:(i + 1) * B].cuda(), Y[i * B:(i + 1) * B].cuda(), B) for i in range(48)]
for i in range(12):
    eng.train_step(crit, dev_batches[i])
torch.cuda.synchronize()
K = 300


def run(name, body):
    torch.cuda.synchronize()
    body(20)
    torch.cuda.synchronize()
    t0 = time.time()
    body(K)
    torch.cuda.synchronize()
    print(f"{name:58s} {(time.time() - t0) / K * 1e6:8.1f} us/step")


def v_dev(n):
    for i in range(n):
        eng.train_step(crit, dev_batches[i % 48])


def v_dev_loss(n):
    eng.async_loss = True
    for i in range(n):
        eng._read_loss(eng.train_step(crit, dev_batches[i % 48]))


def v_dev_loss_sync(n):
    for i in range(n):
        float(eng.train_step(crit, dev_batches[i % 48]))


def v_prefetch(n):
    it = Prefetcher(fs.batches(B, train=True, drop_last=True, epochs=1000), ctx.device)
    for i, b in zip(range(n), it):
        eng.train_step(crit, b)


def v_prefetch_loss(n):
    eng.async_loss = True
    it = Prefetcher(fs.batches(B, train=True, drop_last=True, epochs=1000), ctx.device)
    for i, b in zip(range(n), it):
        eng._read_loss(eng.train_step(crit, b))


def v_direct(n):
    g = eng._graph
    eng.async_loss = True
    for i in range(n):
        lo = (i % 48) * B
        g["static"][0].copy_(fs.features[0].narrow(0, lo, B), non_blocking=True)
        g["static"][1].copy_(fs.labels[0].narrow(0, lo, B), non_blocking=True)
        eng.fopt.pre_graph_replay()
        g["graph"].replay()
        eng._read_loss(g["loss"])


def v_fit(n):
    eng.async_loss = True
    est.fit(fs, batch_size=B, max_steps=eng.state["neval"] - 1 + n, log_every=1, shard=False)


run("device-resident batches, no loss read", v_dev)
run("device-resident batches, lag-1 loss read", v_dev_loss)
run("device-resident batches, synchronous loss read", v_dev_loss_sync)
run("prefetcher (pinned H2D on side stream), no loss read", v_prefetch)
run("prefetcher + lag-1 loss read", v_prefetch_loss)
run("H2D straight into the graph's static inputs + lag-1 loss", v_direct)
run("Estimator.fit (public API)", v_fit)
