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
for mode in ("nccl", "p2p", "nccl-adam", "p2p-owner", "nccl-clip", "p2p-clip"):
    m = copy.deepcopy(base)
    # "p2p": small arenas take the one-handshake push all-reduce with a bf16 wire (csrc/comm/ar_push.cu);
    # "p2p-owner": force the owner-computes reduce-scatter + optimizer + all-gather kernel (csrc/comm/rs_opt_ag.cu, fp32 wire)
    os.environ["ZOO_K1_SMALL"] = "0" if mode == "p2p-owner" else "auto"
    # the owner-computes path moves fp32 and must track NCCL + Adam to rounding; the push path rounds gradients to bf16, which
    # Adam's normalisation would turn into +-lr steps on near-zero gradients, so wire-format parity is checked with SGD (linear)
    opt = torch.optim.Adam(m.parameters(), lr=1e-2) if mode in ("nccl-adam", "p2p-owner") else torch.optim.SGD(m.parameters(), lr=0.05)
    eng = TrainingEngine(m, opt, ctx=ctx, grad_sync="nccl" if mode.startswith("nccl") else "p2p")
    eng.use_cuda_graph = False
    if mode.endswith("clip"):
        eng.set_gradient_clipping_by_l2_norm(0.2)
    eng.prepare()
    assert eng._dp.name == ("allreduce" if mode.startswith("nccl") else "p2p"), eng._dp.name
    if mode.startswith("p2p"):
        assert eng._dp.small == (mode != "p2p-owner")
        assert eng._dp.graph_safe() or mode == "p2p-owner"
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
    eng.close()
d_owner = (results["nccl-adam"] - results["p2p-owner"]).abs().max().item()
d_push = (results["nccl"] - results["p2p"]).abs().max().item()        # gradients travel as bf16: Adam steps differ at rounding size
d_clip = (results["nccl-clip"] - results["p2p-clip"]).abs().max().item()
moved = (results["nccl-clip"] - results["nccl"]).abs().max().item()     # the clip must actually bite
assert d_owner < 1e-4 and d_push < 2e-3 and d_clip < 2e-3 and moved > 2e-3, (d_owner, d_push, d_clip, moved)
torch.distributed.barrier()
if ctx.rank == 0:
    print(f"MGPU_DP_OK world={ctx.world_size} max|nccl-owner|={d_owner:.2e} max|nccl-push|={d_push:.2e} max|clip nccl-push|={d_clip:.2e}")
