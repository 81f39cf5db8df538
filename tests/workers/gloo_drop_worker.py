"""torchrun worker (CPU, gloo, world_size=2): gradient dropping (zoo.pipeline.straggler; reference knobs Topology.scala:1096-1100).
 1. With no drop the micro-batched step equals the one-process step on the whole batch.
 2. Rank 1 is made slow: after the warm-up window it drops micro-batches, the step is normalised by the samples really processed
    (checked against a hand-computed SGD step), both ranks keep identical weights.
 3. With maxDropPercentage below what is dropped the iteration is discarded on both ranks."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ["ZOO_FORCE_CPU"] = "1"
import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from zoo.common.nncontext import init_nncontext
from zoo.common.triggers import MaxIteration
from zoo.feature.common import FeatureSet
from zoo.pipeline.api.keras.optimizers import SGD
from zoo.pipeline.estimator import Estimator

ctx = init_nncontext()
assert ctx.world_size == 2 and ctx.backend == "gloo", ctx
W, r = ctx.world_size, ctx.rank


def mlp():
    torch.manual_seed(5)
    return nn.Sequential(nn.Linear(6, 8), nn.Tanh(), nn.Linear(8, 1))


rng = np.random.RandomState(1)
x = rng.randn(64, 6).astype(np.float32)
y = (x @ rng.randn(6, 1)).astype(np.float32)
half = slice(r * 32, (r + 1) * 32)


def same_on_all_ranks(model):
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    both = [torch.empty_like(flat) for _ in range(W)]
    dist.all_gather(both, flat)
    assert torch.equal(both[0], both[1])


def sgd_reference(rows, lr=0.1):
    ref = mlp()
    xb, yb = torch.from_numpy(x[rows]), torch.from_numpy(y[rows])
    nn.functional.mse_loss(ref(xb), yb).backward()
    with torch.no_grad():
        for p in ref.parameters():
            p.sub_(lr * p.grad)
    return ref


# ---- 1. nothing dropped: equals the whole-batch step ------------------------------------------------------------------------
m = mlp()
est = Estimator(m, SGD(learningrate=0.1)).set_drop_module_property(0.3, 0.5, batch_size=2, warmup_iteration=100)
est.engine.train(FeatureSet.numpy(x[half], y[half], shuffle=False), "mse", MaxIteration(1), batch_size=64, shard=False)
ref = sgd_reference(np.arange(64))
for a, b in zip(m.parameters(), ref.parameters()):
    assert torch.allclose(a, b, atol=1e-6), (a - b).abs().max()
assert est.engine._dropper.stats["dropped_micro_batches"] == 0
same_on_all_ranks(m)

# ---- 2. rank 1 is slow: it drops, the step divides by what was processed ----------------------------------------------------
m = mlp()
est = Estimator(m, SGD(learningrate=0.1)).set_drop_module_property(0.75, 0.8, batch_size=2, warmup_iteration=1)
d = est.engine._dropper
# sleeps dominate the step so that host jitter cannot decide a drop; the fast rank is a little faster after the window than the
# threshold taken from it (an entry AT the threshold is borderline by construction)
d.delay = lambda rank, j: 0.2 if rank == 1 else (0.03 if d.iteration <= 3 else 0.02)
fs = FeatureSet.numpy(x[half], y[half], shuffle=False)
est.engine.train(fs, "mse", MaxIteration(3), batch_size=64, shard=False)          # warm-up, then the window: 2 iterations x 2 ranks
assert d.threshold is not None and d.threshold < 0.1, d.threshold                  # third largest of 4 entries: the fast rank's
assert d.stats["dropped_micro_batches"] == 0
before = [p.detach().clone() for p in m.parameters()]
est.engine.train(fs, "mse", MaxIteration(4), batch_size=64, shard=False)          # iteration 4 runs against the threshold
assert d.stats["dropped_micro_batches"] == 3, d.stats                               # rank 1 ran its first micro-batch only
rows = np.concatenate([np.arange(0, 32), np.arange(32, 40)])                      # rank 0: all 32; rank 1: first 8 of its 32
ref = mlp()
with torch.no_grad():
    for p, b in zip(ref.parameters(), before):
        p.copy_(b)
nn.functional.mse_loss(ref(torch.from_numpy(x[rows])), torch.from_numpy(y[rows])).backward()
for a, b, p in zip(m.parameters(), before, ref.parameters()):
    assert torch.allclose(a, b - 0.1 * p.grad, atol=1e-6), (a - (b - 0.1 * p.grad)).abs().max()
same_on_all_ranks(m)

# ---- 3. more dropped than maxDropPercentage: the iteration is discarded -----------------------------------------------------
d.max_drop_percentage = 0.25                                                        # 3 of 8 dropped > 2 allowed
before = [p.detach().clone() for p in m.parameters()]
est.engine.train(fs, "mse", MaxIteration(5), batch_size=64, shard=False)
assert d.stats["discarded_iterations"] == 1, d.stats
for a, b in zip(m.parameters(), before):
    assert torch.equal(a, b)
same_on_all_ranks(m)

dist.barrier()
if r == 0:
    print("GLOO_DROP_OK")
