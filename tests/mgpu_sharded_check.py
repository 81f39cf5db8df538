"""torchrun worker (N GPUs): NeuralCF with row-sharded tables -- routed ids, one-sided NVLink gathers, owner-side grouped sparse
optimizer, push all-reduce of the dense part -- must match

  (a) a plain-PyTorch oracle on the concatenated global batch (fp32 `nn.Embedding(sparse=True)` + `SparseAdam` for the tables,
      `Adam` for the towers -- none of this repository's kernels), and
  (b) this repository's own single-GPU engine on the concatenated batch,

in eager mode AND when the step is replayed from a CUDA graph (the path every multi-GPU benchmark number uses), with uniform and
with Zipf(1.05) ids (heavy duplicates, skewed owners).  Then a checkpoint round trip: save (one shard file per rank), scramble,
load, compare.  Reference for the test shape: DistriEstimatorSpec.scala:116-206 (multi-partition training equals local training).
"""
import os
import shutil
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn as nn

from zoo.common.nncontext import ZooContext, init_nncontext
from zoo.feature.common import MiniBatch
from zoo.models.recommendation import NeuralCF
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.engine import TrainingEngine

ctx = init_nncontext()
W, r = ctx.world_size, ctx.rank
U, I, C, B, STEPS = 997, 503, 5, 512, 8
dims = dict(user_embed=32, item_embed=32, hidden_layers=[64, 32], mf_embed=16)
crit = SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None)
LR = 1e-2
SGD_LR = 0.5


def zipf(n, rows, g, alpha=1.05):
    u = torch.rand(n, generator=g, dtype=torch.float64)
    a = 1.0 - alpha
    return (((u * ((rows + 1.0) ** a - 1.0)) + 1.0) ** (1.0 / a)).floor().clamp_(1, rows).long()


def batches(kind):
    g = torch.Generator().manual_seed(7)
    if kind == "zipf":
        xs = [torch.stack([zipf(W * B, U, g), zipf(W * B, I, g)], 1) for _ in range(STEPS)]
    else:
        xs = [torch.stack([torch.randint(1, U + 1, (W * B,), generator=g), torch.randint(1, I + 1, (W * B,), generator=g)], 1) for _ in range(STEPS)]
    ys = [torch.randint(1, C + 1, (W * B,), generator=g) for _ in range(STEPS)]
    return xs, ys


class TorchNCF(nn.Module):
    """The oracle: the NeuralCF graph (NeuralCF.scala:45-103) in plain fp32 PyTorch."""

    def __init__(self, src):
        super().__init__()
        tabs = {n: m for n, m in src.named_modules() if hasattr(m, "rows_local")}
        self.emb = nn.ModuleDict({n: nn.Embedding.from_pretrained(t.weight.detach().clone().float(), freeze=False, sparse=True) for n, t in tabs.items()})
        self.lin = nn.ModuleList()
        for layer in list(src.mlp) + [src.linear_last]:
            l = nn.Linear(layer.in_features, layer.out_features)
            l.weight.data.copy_(layer.weight.detach())
            l.bias.data.copy_(layer.bias.detach())
            self.lin.append(l)

    def forward(self, x):
        u, i = x[:, 0], x[:, 1]
        h = torch.cat([self.emb["mlp_user_embed"](u), self.emb["mlp_item_embed"](i)], 1)
        for l in self.lin[:-1]:
            h = torch.relu(l(h))
        return self.lin[-1](torch.cat([h, self.emb["mf_user_embed"](u) * self.emb["mf_item_embed"](i)], 1))


def run_case(kind, use_graph, opt_name="adam"):
    xs, ys = batches(kind)
    mk = (lambda ps: torch.optim.Adam(ps, lr=LR)) if opt_name == "adam" else (lambda ps: torch.optim.SGD(ps, lr=SGD_LR))
    torch.manual_seed(0)
    ref = NeuralCF(U, I, C, sparse=True, **dims)                      # full tables, CPU init: the common starting point
    full = {n: t.weight.detach().clone() for n, t in ref.named_modules() if hasattr(t, "rows_local")}
    dense0 = {n: p.detach().clone() for n, p in ref.named_parameters() if "embed" not in n}
    torch.manual_seed(0)
    sh = NeuralCF(U, I, C, sparse=True, world=W, rank=r, device=ctx.device, **dims)
    with torch.no_grad():                                              # shard row k of rank r = global row k * W + r
        for n, t in sh.named_modules():
            if hasattr(t, "rows_local"):
                rows = full[n][r::W].to(ctx.device)
                t.weight[: rows.shape[0]].copy_(rows)
                t.refresh_shadow()
        for n, p in sh.named_parameters():
            if n in dense0:
                p.copy_(dense0[n])
    torch.cuda.synchronize()
    dist.barrier()
    eng = TrainingEngine(sh, mk([p for p in sh.parameters() if p.requires_grad]), ctx=ctx, grad_sync="p2p")
    eng.use_cuda_graph = use_graph
    eng.prepare()
    assert eng._dp.name == "p2p"
    losses = []
    for s in range(STEPS):
        x, y = xs[s][r * B:(r + 1) * B].cuda(), ys[s][r * B:(r + 1) * B].cuda()
        losses.append(eng.train_step(crit.cuda(), MiniBatch(x, y, B)).detach().clone())
    torch.cuda.synchronize()
    if use_graph:
        assert eng._graph is not None and eng._graph["graph"] is not None, "the step was not replayed from a CUDA graph"
    lt = torch.stack(losses)
    dist.all_reduce(lt)
    lt /= W
    ok, msg = True, ""
    if r == 0:
        # (a) plain-PyTorch oracle
        orc = TorchNCF(ref).cuda()
        so = torch.optim.SparseAdam(list(orc.emb.parameters()), lr=LR) if opt_name == "adam" else torch.optim.SGD(list(orc.emb.parameters()), lr=SGD_LR)
        do = mk(list(orc.lin.parameters()))
        ol = []
        for s in range(STEPS):
            so.zero_grad(); do.zero_grad()
            loss = nn.functional.cross_entropy(orc(xs[s].cuda()), ys[s].cuda() - 1)
            loss.backward()
            so.step(); do.step()
            ol.append(loss.detach())
        ol = torch.stack(ol)
        d_orc = (ol - lt).abs().max().item()
        t_orc = 0.0
        for n, t in sh.named_modules():
            if hasattr(t, "rows_local"):
                want = orc.emb[n].weight.detach()[r::W]
                t_orc = max(t_orc, (want - t.weight.detach()[: want.shape[0]]).abs().mean().item())
        # (b) this repository's single-GPU engine
        solo = ZooContext(0, 0, 1, ctx.device, None, ctx.conf)
        reng = TrainingEngine(ref, mk([p for p in ref.parameters() if p.requires_grad]), ctx=solo)
        reng.use_cuda_graph = False
        reng.prepare()
        rl = torch.stack([reng.train_step(crit.cuda(), MiniBatch(xs[s].cuda(), ys[s].cuda(), W * B)).detach().clone() for s in range(STEPS)])
        torch.cuda.synchronize()
        dl = (rl - lt).abs().max().item()
        worst, mean = 0.0, 0.0
        for n, t in sh.named_modules():
            if hasattr(t, "rows_local"):
                want = dict(ref.named_modules())[n].weight.detach()[r::W]
                got = t.weight.detach()[: want.shape[0]]
                worst = max(worst, (want - got).abs().max().item())
                mean = max(mean, (want - got).abs().mean().item())
        dw = max((p.detach() - dict(ref.named_parameters())[n].detach()).abs().max().item() for n, p in sh.named_parameters() if n in dense0)
        msg = (f"[{kind} graph={use_graph}] vs torch oracle: loss {d_orc:.2e} table mean {t_orc:.2e} | vs 1-GPU engine: loss {dl:.2e} "
               f"table max {worst:.2e} mean {mean:.2e} dense {dw:.2e}")
        print(msg, flush=True)
        msg = f"[{opt_name}]" + msg
        if opt_name == "adam":
            # Adam turns rounding noise on near-zero gradients into +-lr steps (and the dense part travels as bf16 when world > 1),
            # so single weights drift by a few lr; a missed, doubled or mis-routed update shows in the loss trajectory, which
            # must agree tightly, and in the mean.  The torch oracle is fp32 end to end, hence its wider band.
            ok = dl < 1e-3 and worst < 8e-2 and mean < 6e-3 and dw < 6e-2 and d_orc < 2e-2 and t_orc < 2e-2
        else:
            # SGD is linear in the gradient: every weight must agree at rounding size
            ok = dl < 1e-3 and worst < 5e-3 and mean < 2e-4 and dw < 5e-3 and d_orc < 2e-2 and t_orc < 1e-3
    flag = torch.tensor([1 if ok else 0], device=ctx.device)
    dist.broadcast(flag, src=0)
    assert int(flag) == 1, msg
    return eng, sh


eng = None
for kind in ("uniform", "zipf"):
    for use_graph in (False, True):
        if eng is not None:
            eng.close()
        eng, sh = run_case(kind, use_graph)
eng.close()
eng, sh = run_case("zipf", True, "sgd")       # linear optimizer: weights of every table must agree at rounding size
eng.close()
eng, sh = run_case("uniform", True, "adam")   # (the checkpoint round trip below wants Adam state)

# ---- checkpoint round trip with sharded tables (one shard file per rank) ---------------------------------------------------------
box = [tempfile.mkdtemp(prefix="zoo-mgpu-ckpt-")] if r == 0 else [None]
dist.broadcast_object_list(box, src=0)
eng.set_checkpoint(box[0])
eng.state["neval"] = STEPS + 1
before = {n: t.weight.detach().clone() for n, t in sh.named_modules() if hasattr(t, "rows_local")}
m_before = {n: t.exp_avg.detach().clone() for n, t in sh.named_modules() if hasattr(t, "rows_local")}
dense_before = eng.fopt.arena.flat_p.detach().clone()
path = eng.save_checkpoint()
assert os.path.exists(f"{path}.shard-{r}-of-{W}"), "rank shard file missing"
with torch.no_grad():
    for n, t in sh.named_modules():
        if hasattr(t, "rows_local"):
            t.weight.add_(1.0)
            t.exp_avg.zero_()
    eng.fopt.arena.flat_p.mul_(0.5)
eng.load_checkpoint(path)
for n, t in sh.named_modules():
    if hasattr(t, "rows_local"):
        assert torch.equal(t.weight.detach(), before[n]) and torch.equal(t.exp_avg, m_before[n]), n
        assert torch.equal(t.weight16.float(), before[n].to(torch.bfloat16).float()), f"{n}: bf16 shadow not refreshed"
assert torch.equal(eng.fopt.arena.flat_p, dense_before)
dist.barrier()
if r == 0:
    shutil.rmtree(box[0], ignore_errors=True)
    print(f"MGPU_SHARDED_OK world={W}")
