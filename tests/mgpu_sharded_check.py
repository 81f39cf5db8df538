"""torchrun worker: NeuralCF with row-sharded tables (one-sided NVLink gather forward, pull-based sparse
optimizer backward) must match a single-GPU run on the concatenated global batch."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from zoo.common.nncontext import init_nncontext
from zoo.feature.common import MiniBatch
from zoo.models.recommendation import NeuralCF
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.engine import TrainingEngine

ctx = init_nncontext()
W, r = ctx.world_size, ctx.rank
U, I, C, B, STEPS = 997, 503, 5, 512, 6
dims = dict(user_embed=32, item_embed=32, hidden_layers=[64, 32], mf_embed=16)
g = torch.Generator().manual_seed(7)
xs = [torch.stack([torch.randint(1, U + 1, (W * B,), generator=g), torch.randint(1, I + 1, (W * B,), generator=g)], 1) for _ in range(STEPS)]
ys = [torch.randint(1, C + 1, (W * B,), generator=g) for _ in range(STEPS)]
crit = SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None)

# reference: one GPU, whole batch, ordinary (world=1) sparse tables
torch.manual_seed(0)
ref = NeuralCF(U, I, C, sparse=True, **dims)
full = {n: t.weight.detach().clone() for n, t in ref.named_modules() if hasattr(t, "rows_local")}
dense0 = {n: p.detach().clone() for n, p in ref.named_parameters() if "embed" not in n}

torch.manual_seed(0)
sh = NeuralCF(U, I, C, sparse=True, world=W, rank=r, device=ctx.device, **dims)
with torch.no_grad():                      # same initial values: shard row k of rank r = global row k*W + r
    for n, t in sh.named_modules():
        if hasattr(t, "rows_local"):
            rows = full[n][r::W].to(ctx.device)
            t.weight[: rows.shape[0]].copy_(rows)
    for n, p in sh.named_parameters():
        if n in dense0:
            p.copy_(dense0[n])
torch.cuda.synchronize()
dist.barrier()

for use_graph in (False, True):
    eng = TrainingEngine(sh, torch.optim.Adam([p for p in sh.parameters() if p.requires_grad], lr=1e-2), ctx=ctx, grad_sync="p2p")
    eng.use_cuda_graph = use_graph
    eng.prepare()
    if use_graph:
        break
    losses = []
    for s in range(STEPS):
        x, y = xs[s][r * B:(r + 1) * B].cuda(), ys[s][r * B:(r + 1) * B].cuda()
        losses.append(eng.train_step(crit.cuda(), MiniBatch(x, y, B)).detach().clone())
    torch.cuda.synchronize()
    lt = torch.stack(losses)
    dist.all_reduce(lt)
    lt /= W
    break

ok = True
if r == 0:
    # single-process reference on rank 0's GPU (LocalSync, no collectives)
    from zoo.common.nncontext import ZooContext

    solo = ZooContext(0, 0, 1, ctx.device, None, ctx.conf)
    reng = TrainingEngine(ref, torch.optim.Adam([p for p in ref.parameters() if p.requires_grad], lr=1e-2), ctx=solo)
    reng.use_cuda_graph = False
    reng.prepare()
    rl = [reng.train_step(crit.cuda(), MiniBatch(xs[s].cuda(), ys[s].cuda(), W * B)).detach().clone() for s in range(STEPS)]
    torch.cuda.synchronize()
    rl = torch.stack(rl)
    dl = (rl - lt).abs().max().item()
    worst, mean = 0.0, 0.0
    for n, t in sh.named_modules():
        if hasattr(t, "rows_local"):
            want = dict(ref.named_modules())[n].weight.detach()[r::W]
            got = t.weight.detach()[: want.shape[0]]
            worst = max(worst, (want - got).abs().max().item())
            mean = max(mean, (want - got).abs().mean().item())
    dw = max((p.detach() - dict(ref.named_parameters())[n].detach()).abs().max().item() for n, p in sh.named_parameters() if n in dense0)
    print(f"loss diff {dl:.2e}, table diff max {worst:.2e} mean {mean:.2e}, dense diff {dw:.2e}")
    # Adam at lr 1e-2 turns rounding noise on near-zero gradients into +-lr steps: single elements may drift by a few lr,
    # a missed or doubled update would show in the loss and in the mean
    ok = dl < 1e-3 and worst < 4e-2 and mean < 1e-3 and dw < 3e-2
flag = torch.tensor([1 if ok else 0], device=ctx.device)
dist.broadcast(flag, src=0)
dist.barrier()
assert int(flag) == 1
if r == 0:
    print(f"MGPU_SHARDED_OK world={W}")
