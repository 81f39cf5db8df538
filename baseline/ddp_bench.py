#!/usr/bin/env python
"""Same-box NCCL baseline: the SAME models and configs as bench.py written the ordinary way -- plain PyTorch modules,
cuBLAS/cuDNN kernels, `torch.optim`, DistributedDataParallel and `all_to_all_single` over NCCL -- timed with the same rules
(CUDA events, max over ranks, W >= 3 warm-up steps, a distinct random batch every step).  This is "the path that only calls
NCCL" the product is measured against; none of zoo's kernels or engine run here.

    python baseline/ddp_bench.py --config ncf|wnd|resnet50 [--steps K --warmup W]
    torchrun --nproc-per-node N baseline/ddp_bench.py --config ncf|wnd|resnet50 ...

`run_baseline(...)` is also called in-process by bench.py (same process group, right before the product arm) so every bench line
carries its own same-box anchor.

NCF (BASELINE.json config 3) at N GPUs is the textbook model-parallel embedding: tables row-sharded `row % N`, ids bucketed by
owner and exchanged with `all_to_all_single`, rows looked up on the owner (`F.embedding(sparse=True)`), pooled rows returned by a
second all-to-all (autograd-aware, so gradient rows travel back the same way), `SparseAdam` on the owner's shard, DDP + Adam for
the towers, bf16 autocast.  Two tunings a competent user would make are included: the MLP and MF tables of one id column are ONE
192-wide table (one exchange per column instead of two), and ids are exchanged once per column.  The reference does the dense
version of this step with BigDL AllReduceParameter (zoo/src/main/scala/.../pipeline/api/keras/models/Topology.scala:1128-1206).
"""
from __future__ import annotations

import argparse
import json
import os
import time

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------ NCF, row-sharded over NCCL --
class _ShardedTable(nn.Module):
    """Rows `r` with `r % W == rank` of a [rows, D] table; `lookup(ids)` returns the rows of THIS rank's ids, fetched from
    their owners with two all-to-alls (ids out, rows back); backward sends gradient rows to the owners the same way."""

    def __init__(self, rows, D, world, rank, dev):
        super().__init__()
        self.W, self.rank = world, rank
        self.emb = nn.Embedding((rows + world - 1) // world, D, sparse=True, device=dev)
        nn.init.uniform_(self.emb.weight, -0.05, 0.05)

    def lookup(self, ids):
        W = self.W
        if W == 1:
            return self.emb(ids)
        import torch.distributed as dist
        from torch.distributed.nn.functional import all_to_all_single

        owner = ids % W
        order = torch.argsort(owner, stable=True)
        send_counts = torch.bincount(owner, minlength=W)
        recv_counts = torch.empty_like(send_counts)
        dist.all_to_all_single(recv_counts, send_counts)
        sc, rc = send_counts.tolist(), recv_counts.tolist()           # split sizes live on the host (the stock API needs them)
        send_ids = (ids // W)[order]
        recv_ids = torch.empty(sum(rc), dtype=ids.dtype, device=ids.device)
        dist.all_to_all_single(recv_ids, send_ids, rc, sc)
        rows = self.emb(recv_ids)                                      # owner-side gather (sparse gradient)
        back = all_to_all_single(torch.empty(sum(sc), rows.shape[1], dtype=rows.dtype, device=rows.device), rows, sc, rc)
        out = torch.empty_like(back)
        out[order] = back
        return out


class ShardedNCF(nn.Module):
    def __init__(self, U, I, world, rank, dev, C=5, ue=128, ie=128, hidden=(256, 128, 64), mf=64):
        super().__init__()
        self.ue, self.ie, self.mf = ue, ie, mf
        self.user = _ShardedTable(U + 1, ue + mf, world, rank, dev)     # MLP-user | MF-user rows side by side
        self.item = _ShardedTable(I + 1, ie + mf, world, rank, dev)
        layers, prev = [], ue + ie
        for h in hidden:
            layers += [nn.Linear(prev, h), nn.ReLU()]
            prev = h
        self.mlp = nn.Sequential(*layers).to(dev)
        self.last = nn.Linear(prev + mf, C).to(dev)

    def tower(self, u, i):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            h = self.mlp(torch.cat([u[:, :self.ue], i[:, :self.ie]], 1))
            return self.last(torch.cat([h, u[:, self.ue:] * i[:, self.ie:]], 1)).float()


class _Tower(nn.Module):
    """The dense part as its own module so DDP wraps only replicated parameters."""

    def __init__(self, ncf):
        super().__init__()
        self.mlp, self.last, self.ue, self.ie = ncf.mlp, ncf.last, ncf.ue, ncf.ie

    def forward(self, u, i):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            h = self.mlp(torch.cat([u[:, :self.ue], i[:, :self.ie]], 1))
            return self.last(torch.cat([h, u[:, self.ue:] * i[:, self.ie:]], 1)).float()


def _zipf_ids(n, rows, gen, alpha):
    """Zipf(alpha)-distributed ids in [1, rows] (inverse-CDF of the continuous power law; heavy duplicates for alpha ~ 1)."""
    u = torch.rand(n, generator=gen, dtype=torch.float64)
    if abs(alpha - 1.0) < 1e-9:
        x = torch.exp(u * np.log(rows + 1.0))
    else:
        a = 1.0 - alpha
        x = ((u * ((rows + 1.0) ** a - 1.0)) + 1.0) ** (1.0 / a)
    return x.floor().clamp_(1, rows).to(torch.int64)


def make_ncf_batch(B, U, I, gen, zipf=0.0):
    if zipf > 0:
        u, i = _zipf_ids(B, U, gen, zipf), _zipf_ids(B, I, gen, zipf)
    else:
        u, i = torch.randint(1, U + 1, (B,), generator=gen), torch.randint(1, I + 1, (B,), generator=gen)
    return torch.stack([u, i], 1), torch.randint(1, 6, (B,), generator=gen)


def _build_ncf(world, rank, dev, batch, scale, zipf, compile_tower=False):
    U, I = int(12_500_000 * scale) * world, int(1_250_000 * scale) * world
    B = batch or 65536
    model = ShardedNCF(U, I, world, rank, dev)
    tower = _Tower(model)
    net = tower
    if world > 1:
        net = nn.parallel.DistributedDataParallel(tower, device_ids=[dev.index], gradient_as_bucket_view=True)
    if compile_tower:
        net = torch.compile(net)
    opts = [torch.optim.SparseAdam(list(model.user.parameters()) + list(model.item.parameters()), lr=1e-3),
            torch.optim.Adam(tower.parameters(), lr=1e-3)]
    crit = nn.CrossEntropyLoss()
    g = torch.Generator().manual_seed(1234 + rank)

    def mk():
        x, y = make_ncf_batch(B, U, I, g, zipf)
        return x, y - 1

    def step(x, y):
        for o in opts:
            o.zero_grad(set_to_none=True)
        u, i = model.user.lookup(x[:, 0]), model.item.lookup(x[:, 1])
        loss = crit(net(u, i), y)
        loss.backward()
        for o in opts:
            o.step()
        return loss

    return dict(B=B, mk=mk, step=step, name="NCF training samples/sec", unit="samples/s", keep=(model, net),
                model=f"NeuralCF({U}x{I} rows, dim 128+64, MLP 256-256-128-64), tables row-sharded x{world} via all_to_all_single")


# ---------------------------------------------------------------------------------------------------- W&D / ResNet-50 --
def _build_wnd(world, rank, dev, batch):
    B = batch or 32768
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
    net = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], gradient_as_bucket_view=True) if world > 1 else model
    opts = [torch.optim.Adagrad(list(model.deep.parameters()) + list(model.occ.parameters()), lr=1e-3), torch.optim.SGD(model.wide.parameters(), lr=5e-3)]
    crit = nn.CrossEntropyLoss()
    rng = np.random.RandomState(rank)

    def mk():
        w = torch.from_numpy(rng.randint(0, wide_dims, (B, 8)))
        i = torch.from_numpy((rng.rand(B, ind) < 0.1).astype(np.float32))
        e = torch.from_numpy(rng.randint(1, 1001, (B, 1)))
        c = torch.from_numpy(rng.rand(B, cont).astype(np.float32))
        return (w, i, e, c), torch.from_numpy(rng.randint(0, 2, B))

    def step(x, y):
        for o in opts:
            o.zero_grad(set_to_none=True)
        loss = crit(net(*x), y)
        loss.backward()
        for o in opts:
            o.step()
        return loss

    return dict(B=B, mk=mk, step=step, name="Wide&Deep training samples/sec", unit="samples/s", keep=(model, net),
                model="WideAndDeep(Census columns, hidden 100-75-50-25), DDP")


def _build_resnet50(world, rank, dev, batch):
    import torchvision

    B = batch or 256
    model = torchvision.models.resnet50(num_classes=1000).to(dev).to(memory_format=torch.channels_last)
    net = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index], gradient_as_bucket_view=True) if world > 1 else model
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    crit = nn.CrossEntropyLoss()
    mean = torch.tensor([123.68, 116.779, 103.939], device=dev).view(1, 3, 1, 1)
    std = torch.tensor([58.393, 57.12, 57.375], device=dev).view(1, 3, 1, 1)
    g = torch.Generator().manual_seed(1234 + rank)
    mk = lambda: (torch.randint(0, 256, (B, 224, 224, 3), generator=g, dtype=torch.uint8), torch.randint(0, 1000, (B,), generator=g))

    def step(x, y):
        opt.zero_grad(set_to_none=True)
        xin = ((x.permute(0, 3, 1, 2).float() - mean) / std).contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = net(xin)
        loss = crit(out.float(), y)
        loss.backward()
        opt.step()
        return loss

    return dict(B=B, mk=mk, step=step, name="ResNet-50 training images/sec", unit="images/s", keep=(model, net),
                model="ResNet-50 (torchvision) channels_last bf16 autocast, DDP")


def _to_dev(b, dev):
    x, y = b
    x = tuple(t.to(dev, non_blocking=True) for t in x) if isinstance(x, tuple) else x.to(dev, non_blocking=True)
    return x, y.to(dev, non_blocking=True)


def _pin(b):
    x, y = b
    return (tuple(t.pin_memory() for t in x) if isinstance(x, tuple) else x.pin_memory()), y.pin_memory()


def run_baseline(config, world, rank, dev, steps, warmup, batch=0, scale=1.0, zipf=0.0, e2e=True, compile_tower=False):
    """Returns the result dict (same fields on every rank).  The process group must already exist when world > 1."""
    import torch.distributed as dist

    torch.manual_seed(0)
    if config in ("ncf", "ncf-zipf"):
        cfg = _build_ncf(world, rank, dev, batch, scale, zipf, compile_tower)
    elif config == "wnd":
        cfg = _build_wnd(world, rank, dev, batch)
    elif config == "resnet50":
        cfg = _build_resnet50(world, rank, dev, batch)
    else:
        raise ValueError(f"no NCCL baseline arm for config {config!r}")
    step, B = cfg["step"], cfg["B"]
    Wm = max(warmup, 3)
    host = [_pin(cfg["mk"]()) for _ in range(min(steps + Wm, 32))]
    pool = [_to_dev(b, dev) for b in host]
    barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)
    for i in range(Wm):
        step(*pool[i % len(pool)])
    torch.cuda.synchronize(); barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        loss = step(*pool[(i + Wm) % len(pool)])
    e1.record()
    barrier(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    out = {"impl": "torch-nccl-baseline", "metric": cfg["name"], "value": steps * B * world / (float(ms) / 1e3), "unit": cfg["unit"],
           "n_gpus": world, "steps": steps, "warmup": Wm, "ms_per_step": float(ms) / steps, "dtype": "bf16 autocast", "data": "synthetic",
           "config": {"model": cfg["model"], "global_batch": B * world, "zipf": zipf or None, "torch_compile": bool(compile_tower)},
           "final_loss": float(loss)}
    if e2e:
        # end to end: every step copies its batch from pinned host memory and reads the loss back to the host
        del pool
        nxt = _to_dev(host[0], dev)
        torch.cuda.synchronize(); barrier()
        t0 = time.time()
        for i in range(steps):
            cur, nxt = nxt, _to_dev(host[(i + 1) % len(host)], dev)
            lv = float(step(*cur))
        torch.cuda.synchronize(); barrier()
        dt = torch.tensor([time.time() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        x0, y0 = host[0]
        bi = sum(t.numel() * t.element_size() for t in (x0 if isinstance(x0, tuple) else (x0,))) + y0.numel() * y0.element_size()
        out["e2e"] = {"value": steps * B * world / float(dt), "unit": cfg["unit"], "ms_per_step": float(dt) * 1e3 / steps,
                      "h2d_bytes_per_step": bi, "d2h_bytes_per_step": 4, "final_loss": lv}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="ncf")
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--zipf", type=float, default=0.0)
    ap.add_argument("--compile", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    out = run_baseline(a.config, world, rank, dev, a.steps, a.warmup, a.batch, a.scale, a.zipf, not a.no_e2e, a.compile)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


if __name__ == "__main__":
    main()
