"""torchrun worker: data-parallel training with the fused P2P exchange must match the NCCL
all-reduce + local optimizer path step by step (same data, same init)."""
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn

from zoo.common.nncontext import init_nncontext
from zoo.feature.common import MiniBatch
from zoo.ops.dense import FusedDense, FusedMLP
from zoo.pipeline.api.keras.objectives import MeanSquaredError
from zoo.pipeline.engine import TrainingEngine

ctx = init_nncontext()
torch.manual_seed(0)
base = nn.Sequential(FusedMLP(64, [128, 64]), FusedDense(64, 8))
results = {}
for mode in ("nccl", "p2p"):
    m = copy.deepcopy(base)
    eng = TrainingEngine(m, torch.optim.Adam(m.parameters(), lr=1e-2), ctx=ctx, grad_sync=mode)
    eng.use_cuda_graph = False
    eng.prepare()
    assert eng._dp.name == ("p2p" if mode == "p2p" else "allreduce"), eng._dp.name
    g = torch.Generator().manual_seed(100 + ctx.rank)
    crit = MeanSquaredError()
    for step in range(6):
        x = torch.randn(256, 64, generator=g).cuda()
        y = torch.randn(256, 8, generator=g).cuda()
        loss = eng.train_step(crit, MiniBatch(x, y, 256))
    torch.cuda.synchronize()
    results[mode] = eng.fopt.arena.flat_p.clone()
    # every rank must hold identical weights
    w = results[mode].clone()
    torch.distributed.broadcast(w, src=0)
    assert torch.equal(w, results[mode]), f"{mode}: ranks diverged"
diff = (results["nccl"] - results["p2p"]).abs().max().item()
assert diff < 1e-4, diff
torch.distributed.barrier()
if ctx.rank == 0:
    print(f"MGPU_DP_OK world={ctx.world_size} max|nccl-p2p|={diff:.2e}")
