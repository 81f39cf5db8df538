"""torchrun worker (2 ranks): embedding gather from a table that lives in the PEER process (CUDA IPC mapping),
random rows, for growing table sizes -- shows whether address translation of the IPC mapping limits random access."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from zoo.common.nncontext import init_nncontext
from zoo.native import lib as nl
from zoo.parallel.symm import get_heap

ctx = init_nncontext()
heap = get_heap(ctx)
lib = nl.cuda_lib()
D, B = 128, 65536 * 4
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
if ctx.rank == 0:
    print('symmetric heap mode:', heap.mode)
for rows in (4_000_000, 25_000_000):
    buf = heap.alloc(rows * D * 4)
    t = buf.tensor(torch.float32, rows * D)
    t.normal_()
    torch.cuda.synchronize(); dist.barrier()
    out = torch.empty(B, D, dtype=torch.bfloat16, device=ctx.device)
    for pattern in ("random",):
        ids = (torch.randint(0, rows, (B, 1), device=ctx.device) if pattern == "random" else torch.arange(B, device=ctx.device).reshape(B, 1)).long()
        for where in ("local", "peer"):
            src = buf.ptrs[ctx.rank if where == "local" else 1 - ctx.rank]
            arr = (C.c_void_p * 1)(src)
            st = torch.cuda.current_stream().cuda_stream
            run = lambda: nl.check(lib.zoo_embedding_gather(ids.data_ptr(), 1, 0, 0, B, 1, 0, arr, rows, None, 0, 1, D, 0,
                                                            out.data_ptr(), D, None, None, st), "g")
            for only in ((0,), (0, 1)):          # one direction, then both ranks at once
                dist.barrier()
                if ctx.rank in only:
                    run(); torch.cuda.synchronize()
                    e0.record()
                    for _ in range(5):
                        run()
                    e1.record(); torch.cuda.synchronize()
                    us = e0.elapsed_time(e1) * 1e3 / 5
                    if ctx.rank == 0:
                        print(f"rows={rows/1e6:5.1f}M {where:5s} {pattern:10s} ranks={only}: {us:8.1f} us {B * D * 4 / us / 1e3:7.1f} GB/s", flush=True)
                dist.barrier()
