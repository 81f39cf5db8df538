"""torchrun worker: the fused gradient exchange (K1: reduce-scatter + optimizer + all-gather over peer memory) on a ResNet-50 sized
parameter arena, against NCCL all-reduce + the local fused optimizer; device time (CUDA events, max over ranks) and the fraction
of the NVLink roofline.
usage: torchrun --nproc-per-node N benchmarks/k1_bench.py [million_params]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn as nn

from zoo.common.nncontext import init_nncontext
from zoo.pipeline.engine import TrainingEngine

ctx = init_nncontext()
W = ctx.world_size
mp = float(sys.argv[1]) if len(sys.argv) > 1 else 25.6
n = int(mp * 1e6) // 4096 * 4096
rows = n // 4096
out = {}
for mode in ("nccl", "p2p"):
    torch.manual_seed(0)
    m = nn.Linear(4096, rows, bias=False).to(ctx.device)
    eng = TrainingEngine(m, torch.optim.Adam(m.parameters(), lr=1e-3), ctx=ctx, grad_sync=mode)
    eng.use_cuda_graph = False
    eng.prepare()
    g = eng.fopt.arena.flat_g
    ts = []
    for it in range(9):
        g.normal_()
        torch.cuda.synchronize()
        torch.distributed.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng._dp.step()
        e1.record()
        torch.cuda.synchronize()
        if it >= 3:
            ts.append(e0.elapsed_time(e1))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device=ctx.device)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    out[mode] = float(t)
    chk = eng.fopt.arena.flat_p.double().sum().reshape(1)
    lo, hi = chk.clone(), chk.clone()
    torch.distributed.all_reduce(lo, op=torch.distributed.ReduceOp.MIN)
    torch.distributed.all_reduce(hi, op=torch.distributed.ReduceOp.MAX)
    assert float(hi - lo) == 0.0, f"{mode}: replicas diverged"
    del eng, m
    torch.cuda.empty_cache()
if ctx.rank == 0:
    # per GPU and direction: (W-1)/W of the fp32 gradients in, (W-1)/W of the fp32 weights + bf16 shadows out
    out_bytes = (W - 1) / W * n * (4 + 2)
    in_bytes = (W - 1) / W * n * 4
    roof_ms = max(out_bytes, in_bytes) / 900e9 * 1e3
    print(f"K1_BENCH world={W} params={n / 1e6:.1f}M  p2p fused {out['p2p']:.3f} ms  nccl all-reduce + optimizer {out['nccl']:.3f} ms  "
          f"NVLink roofline {roof_ms:.3f} ms -> {100 * roof_ms / out['p2p']:.0f} % of roofline; speed-up {out['nccl'] / out['p2p']:.2f}x")
torch.distributed.barrier()
