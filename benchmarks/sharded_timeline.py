"""torchrun worker: kernel timeline of ONE CUDA-graph replay of the sharded NCF step on rank 0 (torch.profiler / CUPTI):
start offset, duration and gap of every kernel, so barrier waits and exposed latencies are visible.
usage: torchrun --nproc-per-node N benchmarks/sharded_timeline.py"""
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

ctx = init_nncontext()
W = ctx.world_size
U, I, B = 12_500_000 * W, 1_250_000 * W, 65536
m = NeuralCF(U, I, 5, user_embed=128, item_embed=128, hidden_layers=[256, 128, 64], mf_embed=64, world=W, rank=ctx.rank, device=ctx.device)
est = Estimator.from_torch(model=m, optimizer=torch.optim.Adam([p for p in m.parameters() if p.requires_grad], lr=1e-3),
                           loss=SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None), ctx=ctx)
est.engine.prepare()
g = torch.Generator().manual_seed(ctx.rank)
batches = [(torch.stack([torch.randint(1, U + 1, (B,), generator=g), torch.randint(1, I + 1, (B,), generator=g)], 1).cuda(),
            torch.randint(1, 6, (B,), generator=g).cuda()) for _ in range(16)]
crit = est.loss.cuda()
for x, y in batches[:10]:
    est.engine.train_step(crit, MiniBatch(x, y, B))
torch.cuda.synchronize()
ctx.barrier()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for x, y in batches[10:]:
        est.engine.train_step(crit, MiniBatch(x, y, B))
    torch.cuda.synchronize()
if ctx.rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time > 0]
    evs.sort(key=lambda e: e.time_range.start)
    names = [e.name for e in evs]
    ends = [i for i, n in enumerate(names) if "rs_opt_ag" in n or "dense_opt_kernel" in n]     # last compute kernel of a step
    s0, s1 = ends[-3] + 1, ends[-2] + 1
    t0 = evs[s0].time_range.start
    prev_end = t0
    print(f"world={W}: one replayed step on rank 0 ({s1 - s0} kernels, {(evs[s1].time_range.start - t0):.1f} us)")
    for e in evs[s0:s1]:
        st, du = e.time_range.start - t0, e.device_time
        print(f"  +{st:8.1f} us  dur {du:7.1f}  gap {st - (prev_end - t0):6.1f}  {e.name[:90]}")
        prev_end = e.time_range.start + du
ctx.barrier()
