"""torchrun worker (CPU, gloo, world_size=2):
 1. gradient clipping acts on the AVERAGED gradient: one clipped SGD step on two half batches equals the one-process step on the
    whole batch (torch.nn.utils.clip_grad_norm_ / clamp on the mean gradient);
 2. distributed retry (Topology.scala:1180-1262): rank 1 fails at iteration 7, every rank reloads the newest checkpoint and
    training completes with identical weights on both ranks."""
import copy
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ["ZOO_FORCE_CPU"] = "1"
import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from zoo.common.nncontext import init_nncontext
from zoo.common.triggers import MaxIteration, SeveralIteration
from zoo.feature.common import FeatureSet
from zoo.pipeline.api.keras.optimizers import SGD, Adam
from zoo.pipeline.estimator import Estimator

ctx = init_nncontext()
assert ctx.world_size == 2 and ctx.backend == "gloo", ctx
W, r = ctx.world_size, ctx.rank


def mlp():
    torch.manual_seed(3)
    return nn.Sequential(nn.Linear(6, 16), nn.ReLU(), nn.Linear(16, 1))


rng = np.random.RandomState(0)
x = rng.randn(64, 6).astype(np.float32)
y = (x @ rng.randn(6, 1) * 3).astype(np.float32)

# ---- 1. clipping on the averaged gradient -------------------------------------------------------------------------------
for kind in ("l2", "const"):
    ref = mlp()
    loss = nn.functional.mse_loss(ref(torch.from_numpy(x)), torch.from_numpy(y))
    loss.backward()
    if kind == "l2":
        total = torch.sqrt(sum((p.grad ** 2).sum() for p in ref.parameters()))
        assert total > 0.5, total                      # the clip below must actually bite
        torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.5)
    else:
        for p in ref.parameters():
            p.grad.clamp_(-0.05, 0.05)
    with torch.no_grad():
        for p in ref.parameters():
            p.sub_(0.1 * p.grad)
    m = mlp()
    est = Estimator(m, SGD(learningrate=0.1))
    if kind == "l2":
        est.set_l2_norm_gradient_clipping(0.5)
    else:
        est.set_constant_gradient_clipping(-0.05, 0.05)
    half = slice(r * 32, (r + 1) * 32)                  # this rank's half of the global batch
    est.engine.train(FeatureSet.numpy(x[half], y[half], shuffle=False), "mse", MaxIteration(1), batch_size=64, shard=False)
    for a, b in zip(ref.parameters(), m.parameters()):
        assert torch.allclose(a, b.detach(), atol=2e-6), (kind, (a - b).abs().max())
dist.barrier()

# ---- 2. distributed retry -------------------------------------------------------------------------------------------------
box = [tempfile.mkdtemp(prefix="zoo-retry-")] if r == 0 else [None]
dist.broadcast_object_list(box, src=0)
xs = rng.randn(1024, 6).astype(np.float32)
ys = (xs @ rng.randn(6, 1)).astype(np.float32)
m = mlp()
w0 = copy.deepcopy(m.state_dict())
est = Estimator(m, Adam(lr=1e-2), model_dir=box[0])
os.environ["ZOO_FAULT_INJECT"] = "1:7"
est.train(FeatureSet.numpy(xs, ys, shuffle=False), "mse", MaxIteration(12), SeveralIteration(3), batch_size=64)
assert "ZOO_FAULT_INJECT" not in os.environ or r != 1, "the fault did not fire on rank 1"
os.environ.pop("ZOO_FAULT_INJECT", None)
assert est.engine.state["neval"] == 13, est.engine.state
flat = est.engine.fopt.arena.flat_p.clone()
other = flat.clone()
dist.broadcast(other, src=0)
assert torch.equal(flat, other), "ranks diverged after the retry"
# an uninterrupted run: 6 iterations, checkpoint, (the retry restarts the pass from that checkpoint) 6 more = the same 12 updates?
# the retried run repeats iterations 7..12 on the data of a fresh pass, so compare against exactly that schedule
m2 = mlp()
m2.load_state_dict(w0)
est2 = Estimator(m2, Adam(lr=1e-2))
est2.train(FeatureSet.numpy(xs, ys, shuffle=False), "mse", MaxIteration(6), batch_size=64)
est2.train(FeatureSet.numpy(xs, ys, shuffle=False), "mse", MaxIteration(12), batch_size=64)
assert torch.allclose(est2.engine.fopt.arena.flat_p, flat, atol=1e-6), (est2.engine.fopt.arena.flat_p - flat).abs().max()
if r == 0:
    print("GLOO_RETRY_CLIP_OK")
dist.barrier()
