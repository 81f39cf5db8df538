#!/usr/bin/env python
"""Same-box baseline: the SAME models and configs as bench.py written the ordinary way -- plain PyTorch modules,
cuBLAS/cuDNN kernels, `torch.optim`, DistributedDataParallel over NCCL -- timed with the same rules
(CUDA events, max over ranks, W warm-up steps).  This is "the path that only calls NCCL" that the product is
measured against; none of zoo's kernels or engine are used here.

    python baseline/ddp_bench.py --config ncf|wnd|resnet50 [--steps K --warmup W]
    torchrun --nproc-per-node N baseline/ddp_bench.py --config wnd|resnet50 ...

NCF: `nn.Embedding(sparse=True)` + `SparseAdam` for the tables (the stock way to train big tables), Adam for the
towers, bf16 autocast.  Only N=1: replicated 100M x 128 tables with Adam state do not fit one GPU, and torch has no
built-in row-sharded embedding, which is exactly what the product adds.
"""
import argparse
import json
import os

import numpy as np
import torch
import torch.nn as nn

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="ncf")
ap.add_argument("--steps", type=int, default=50)
ap.add_argument("--warmup", type=int, default=5)
ap.add_argument("--batch", type=int, default=0)
a = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(0)
g = torch.Generator().manual_seed(1234 + rank)


class NCF(nn.Module):
    def __init__(self, U, I, C=5, ue=128, ie=128, hidden=(256, 128, 64), mf=64):
        super().__init__()
        self.mu, self.mi = nn.Embedding(U + 1, ue, sparse=True), nn.Embedding(I + 1, ie, sparse=True)
        self.gu, self.gi = nn.Embedding(U + 1, mf, sparse=True), nn.Embedding(I + 1, mf, sparse=True)
        for e in (self.mu, self.mi, self.gu, self.gi):
            nn.init.uniform_(e.weight, -0.05, 0.05)
        layers, prev = [], ue + ie
        for h in hidden:
            layers += [nn.Linear(prev, h), nn.ReLU()]
            prev = h
        self.mlp = nn.Sequential(*layers)
        self.last = nn.Linear(prev + mf, C)

    def forward(self, x):
        u, i = x[:, 0], x[:, 1]
        with torch.autocast("cuda", dtype=torch.bfloat16):
            h = self.mlp(torch.cat([self.mu(u), self.mi(i)], 1))
            return self.last(torch.cat([h, self.gu(u) * self.gi(i)], 1)).float()


if a.config == "ncf":
    assert world == 1, "NCF baseline is single-GPU (see the module docstring)"
    U, I, B = 12_500_000, 1_250_000, a.batch or 65536
    model = NCF(U, I).to(dev)
    sparse_p = [p for n, p in model.named_parameters() if n.split(".")[0] in ("mu", "mi", "gu", "gi")]
    dense_p = [p for n, p in model.named_parameters() if n.split(".")[0] not in ("mu", "mi", "gu", "gi")]
    opts = [torch.optim.SparseAdam(sparse_p, lr=1e-3), torch.optim.Adam(dense_p, lr=1e-3)]
    crit = nn.CrossEntropyLoss()
    mk = lambda: (torch.stack([torch.randint(1, U + 1, (B,), generator=g), torch.randint(1, I + 1, (B,), generator=g)], 1).to(dev),
                  torch.randint(0, 5, (B,), generator=g).to(dev))
    name, unit = "NCF training samples/sec", "samples/s"
elif a.config == "wnd":
    B = a.batch or 32768
    wide_dims, ind, cont = 3049, 38, 5

    class WND(nn.Module):
        def __init__(self):
            super().__init__()
            self.wide = nn.EmbeddingBag(wide_dims, 2, mode="sum")
            self.occ = nn.Embedding(1001, 8)
            dims = [ind + 8 + cont, 100, 75, 50, 25]
            self.deep = nn.Sequential(*[m for a_, b_ in zip(dims[:-1], dims[1:]) for m in (nn.Linear(a_, b_), nn.ReLU())], nn.Linear(25, 2), nn.ReLU())

        def forward(self, w, i, e, c):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                d = self.deep(torch.cat([i, self.occ(e[:, 0]), c], 1))
            return self.wide(w) + d.float()

    model = WND().to(dev)
    opts = [torch.optim.Adagrad(list(model.deep.parameters()) + list(model.occ.parameters()), lr=1e-3), torch.optim.SGD(model.wide.parameters(), lr=5e-3)]
    crit = nn.CrossEntropyLoss()
    rng = np.random.RandomState(rank)

    def mk():
        w = torch.from_numpy(rng.randint(0, wide_dims, (B, 8))).to(dev)
        i = torch.from_numpy((rng.rand(B, ind) < 0.1).astype(np.float32)).to(dev)
        e = torch.from_numpy(rng.randint(1, 1001, (B, 1))).to(dev)
        c = torch.from_numpy(rng.rand(B, cont).astype(np.float32)).to(dev)
        return (w, i, e, c), torch.from_numpy(rng.randint(0, 2, B)).to(dev)

    name, unit = "Wide&Deep training samples/sec", "samples/s"
else:
    import torchvision

    B = a.batch or 256
    model = torchvision.models.resnet50(num_classes=1000).to(dev).to(memory_format=torch.channels_last)
    opts = [torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)]
    crit = nn.CrossEntropyLoss()
    mean, std = torch.tensor([123.68, 116.779, 103.939], device=dev).view(1, 3, 1, 1), torch.tensor([58.393, 57.12, 57.375], device=dev).view(1, 3, 1, 1)
    mk = lambda: (torch.randint(0, 256, (B, 224, 224, 3), generator=g, dtype=torch.uint8).to(dev), torch.randint(0, 1000, (B,), generator=g).to(dev))
    name, unit = "ResNet-50 training images/sec", "images/s"

net = model
if world > 1:
    net = nn.parallel.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True)


def step(x, y):
    for o in opts:
        o.zero_grad(set_to_none=True)
    if a.config == "resnet50":
        xin = ((x.permute(0, 3, 1, 2).float() - mean) / std).contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = net(xin)
        loss = crit(out.float(), y)
    else:
        out = net(*x) if isinstance(x, tuple) else net(x)
        loss = crit(out, y)
    loss.backward()
    for o in opts:
        o.step()
    return loss


pool = [mk() for _ in range(min(a.steps + a.warmup, 32))]
for i in range(max(a.warmup, 3)):
    step(*pool[i % len(pool)])
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(a.steps):
    loss = step(*pool[(i + 3) % len(pool)])
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"impl": "torch-ddp-baseline", "metric": name, "value": a.steps * B * world / (float(ms) / 1e3), "unit": unit, "n_gpus": world,
                      "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": float(ms) / a.steps, "dtype": "bf16 autocast", "data": "synthetic",
                      "config": {"global_batch": B * world}, "final_loss": float(loss)}))
if world > 1:
    dist.destroy_process_group()
