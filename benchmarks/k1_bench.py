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
nvls = {}
for mode in ("nccl", "p2p", "p2p-nonvls"):
    os.environ["ZOO_NVLS"] = "0" if mode == "p2p-nonvls" else "1"      # force both variants: the default (`auto`) picks by world size
    if mode == "p2p-nonvls":
        from zoo.parallel import symm as _symm

        _symm._HEAP._mc_gran = None         # re-decide multicast for the buffers of this mode
    torch.manual_seed(0)
    m = nn.Linear(4096, rows, bias=False).to(ctx.device)
    eng = TrainingEngine(m, torch.optim.Adam(m.parameters(), lr=1e-3), ctx=ctx, grad_sync="nccl" if mode == "nccl" else "p2p")
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
    nvls[mode] = bool(getattr(eng._dp, "nvls", False))
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
    # roofline per the profiling recipe: bytes that must cross NVLink per GPU and direction / 770 GB/s (measured peer copy)
    roof770 = max(out_bytes, in_bytes) / 770e9 * 1e3
    print(f"K1_BENCH world={W} params={n / 1e6:.1f}M  fused (NVLS={nvls['p2p']}) {out['p2p']:.3f} ms  fused peer-pointer path {out['p2p-nonvls']:.3f} ms  "
          f"nccl all-reduce + optimizer {out['nccl']:.3f} ms  NVLink roofline {roof_ms:.3f} ms @900 / {roof770:.3f} ms @770 GB/s -> "
          f"{100 * roof770 / min(out['p2p'], out['p2p-nonvls']):.0f} % of the 770 GB/s roofline; speed-up over nccl {out['nccl'] / min(out['p2p'],
                  out['p2p-nonvls']):.2f}x")
torch.distributed.barrier()
