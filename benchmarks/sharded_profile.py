"""torchrun worker: per-kernel device time of the sharded NCF step (torch.profiler / CUPTI), rank 0 prints the table.
usage: torchrun --nproc-per-node N benchmarks/sharded_profile.py [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from zoo.common.nncontext import init_nncontext
from zoo.feature.common import MiniBatch
from zoo.models.recommendation import NeuralCF
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.estimator import Estimator

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
ctx = init_nncontext()
W = ctx.world_size
U, I, B = 12_500_000 * W, 1_250_000 * W, 65536
m = NeuralCF(U, I, 5, user_embed=128, item_embed=128, hidden_layers=[256, 128, 64], mf_embed=64, world=W, rank=ctx.rank, device=ctx.device)
est = Estimator.from_torch(model=m, optimizer=torch.optim.Adam([p for p in m.parameters() if p.requires_grad], lr=1e-3),
                           loss=SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None), ctx=ctx)
est.engine.use_cuda_graph = False
est.engine.prepare()
g = torch.Generator().manual_seed(ctx.rank)
batches = [(torch.stack([torch.randint(1, U + 1, (B,), generator=g), torch.randint(1, I + 1, (B,), generator=g)], 1).cuda(),
            torch.randint(1, 6, (B,), generator=g).cuda()) for _ in range(steps + 3)]
crit = est.loss.cuda()
for x, y in batches[:3]:
    est.engine.train_step(crit, MiniBatch(x, y, B))
torch.cuda.synchronize()
ctx.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for x, y in batches[3:]:
        est.engine.train_step(crit, MiniBatch(x, y, B))
    torch.cuda.synchronize()
if ctx.rank == 0:
    rows = [(e.key, e.device_time_total / steps, e.count / steps) for e in prof.key_averages() if e.device_time_total > 0]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    print(f"world={W} per-step device time {tot:.1f} us")
    for k, t, c in rows[:25]:
        print(f"{t:9.1f} us  x{c:4.1f}  {k[:110]}")
ctx.barrier()
