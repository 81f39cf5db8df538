"""torchrun worker (CPU, gloo, world_size=2): the NCCL baseline's row-sharded table (baseline/ddp_bench.py) returns the same rows
and produces the same gradient rows as a plain full-table lookup."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from baseline.ddp_bench import _ShardedTable, _zipf_ids

dist.init_process_group("gloo")
W, r = dist.get_world_size(), dist.get_rank()
rows, D, B = 101, 8, 64
torch.manual_seed(0)
full = torch.randn(rows, D)
t = _ShardedTable(rows, D, W, r, torch.device("cpu"))
with torch.no_grad():
    sh = full[r::W]
    t.emb.weight[: sh.shape[0]].copy_(sh)
g = torch.Generator().manual_seed(5 + r)
ids = _zipf_ids(B, rows - 1, g, 1.05)
out = t.lookup(ids)
assert torch.equal(out, full[ids]), (out - full[ids]).abs().max()
coef = torch.randn(B, D, generator=g)
(out * coef).sum().backward()
grad = t.emb.weight.grad.to_dense()
# expected: every rank's contributions to the rows this rank owns
allids, allcoef = [torch.empty_like(ids) for _ in range(W)], [torch.empty_like(coef) for _ in range(W)]
dist.all_gather(allids, ids)
dist.all_gather(allcoef, coef)
want = torch.zeros(rows, D)
for i_, c_ in zip(allids, allcoef):
    want.index_add_(0, i_, c_)
want = want[r::W]
assert torch.allclose(grad[: want.shape[0]], want, atol=1e-5)
dist.barrier()
if r == 0:
    print("GLOO_BASELINE_OK")
