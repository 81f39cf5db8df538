"""Extra bench.py configs (BASELINE.json configs 2, 4, 5): `--config wnd | resnet50 | serving-ncf | serving-resnet50`.

wnd       Wide&Deep on synthetic Census-shaped rows (column info of CensusWideAndDeep.scala:94-109, hidden
          100-75-50-25, 2 classes), FTRL on the wide part + Adagrad on the deep part (per-sub-module optimizers,
          CensusWideAndDeep.scala:122-132), data-parallel with the fused gradient exchange (K1).
resnet50  torchvision ResNet-50, bf16 autocast + channels_last, SGD momentum, 256 images per GPU, data-parallel (K1);
          end to end the batch arrives as uint8 NHWC and is normalised / laid out on the GPU.
serving-* Cluster Serving records/s and p50/p99 latency: enqueue -> broker -> engine (K13 / model / K14) -> results.
"""
from __future__ import annotations

import json
import os
import time

import numpy as np


def build_other(args, ctx, torch):
    if args.config == "wnd":
        return _wnd(args, ctx, torch)
    if args.config == "resnet50":
        return _resnet50(args, ctx, torch)
    if args.config.startswith("serving"):
        return {"custom_run": lambda a, c: _serving(a, c, torch)}
    raise SystemExit(f"unknown --config {args.config}")


def _wnd(args, ctx, torch):
    from zoo.models.recommendation import ColumnFeatureInfo, WideAndDeep
    from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
    from zoo.pipeline.api.keras.optimizers import Adagrad, Ftrl

    info = ColumnFeatureInfo(wide_base_cols=["edu", "mari", "rela", "work", "occ", "age_bucket"], wide_base_dims=[16, 7, 6, 9, 1000, 11],
                             wide_cross_cols=["edu_occ", "age_edu_occ"], wide_cross_dims=[1000, 1000],
                             indicator_cols=["work", "edu", "mari", "rela"], indicator_dims=[9, 16, 7, 6], embed_cols=["occ"],
                             embed_in_dims=[1000], embed_out_dims=[8],
                             continuous_cols=["age", "education_num", "capital_gain", "capital_loss", "hours_per_week"])
    batch = args.batch or 32768
    model = WideAndDeep(2, info, model_type="wide_n_deep", hidden_layers=[100, 75, 50, 25], sparse=False)
    loss = SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None)

    def make_batches(n, device, pin=False):
        rng = np.random.RandomState(1234 + ctx.rank)
        out = []
        for _ in range(n):
            xs = model.example_inputs(batch, rng)
            y = torch.from_numpy(rng.randint(1, 3, batch).astype(np.int64))
            out.append((tuple(t.to(device) for t in xs), y.to(device)) if device.type == "cuda" else (xs, y))
        return out

    W = ctx.world_size
    return dict(model=model, loss=loss, batch=batch, make_batches=make_batches, model_name="WideAndDeep(Census columns, hidden 100-75-50-25)",
                parallelism="single" if W == 1 else f"dp{W} (fused reduce-scatter+optimizer+all-gather)", metric="Wide&Deep training samples/sec",
                optimizer={"wide_part": Ftrl(learningrate=min(5e-3, 1 / np.sqrt(3049))), "deep_part": Adagrad(learningrate=1e-3)},
                l2_policy="a distinct random batch every step; activations 32768 x 100 x bf16 per layer, L2 not flushed between steps",
                extra={"optimizer": "ftrl(wide)+adagrad(deep)", "wide_dims": 3049})


def _resnet50(args, ctx, torch):
    import torch.nn as nn
    import torchvision

    batch = args.batch or 256
    model = torchvision.models.resnet50(num_classes=1000)
    loss = nn.CrossEntropyLoss()

    class UInt8Front(nn.Module):
        """uint8 NHWC batch -> normalised NCHW bf16 on the GPU (one pass), then the network."""

        def __init__(self, net):
            super().__init__()
            self.net = net
            self.register_buffer("mean", torch.tensor([123.68, 116.779, 103.939]).view(1, 3, 1, 1))
            self.register_buffer("std", torch.tensor([58.393, 57.12, 57.375]).view(1, 3, 1, 1))

        def forward(self, x):
            if x.dtype == torch.uint8:
                x = (x.permute(0, 3, 1, 2).float() - self.mean) / self.std
            return self.net(x)

    wrapped = UInt8Front(model)

    def make_batches(n, device, pin=False):
        g = torch.Generator().manual_seed(1234 + ctx.rank)
        out = []
        for _ in range(n):
            x = torch.randint(0, 256, (batch, 224, 224, 3), generator=g, dtype=torch.uint8)
            y = torch.randint(0, 1000, (batch,), generator=g)
            out.append((x.to(device), y.to(device)) if device.type == "cuda" else (x, y))
        return out

    W = ctx.world_size
    opt = torch.optim.SGD(wrapped.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    return dict(model=wrapped, loss=loss, batch=batch, make_batches=make_batches, model_name="ResNet-50 (torchvision), 224x224",
                parallelism="single" if W == 1 else f"dp{W} (fused reduce-scatter+optimizer+all-gather)", metric="ResNet-50 training images/sec",
                unit="images/s", optimizer=opt, engine_opts={"autocast_dtype": torch.bfloat16, "channels_last": True}, e2e_pool=8,
                l2_policy="inputs 256x224x224x3 (38.5 MB uint8) and activations exceed L2; a distinct batch every step",
                extra={"optimizer": "sgd(momentum 0.9, wd 1e-4)", "conv_body": "cuDNN via torch (SURVEY 2.6: conv body stays library)"})


def _serving(args, ctx, torch):
    """Each rank runs one serving engine with its in-process broker (replica parallelism); a feeder thread enqueues
    records continuously; records/s is counted over K batches after W warm-up batches; latency = enqueue -> result written."""
    import threading

    from zoo.serving import ClusterServing, InputQueue

    kind = "resnet50" if "resnet" in args.config else "ncf"
    bs = args.batch or (64 if kind == "resnet50" else 4096)
    dev = ctx.device
    base = None
    if not getattr(args, "no_baseline", False):        # same-box anchor: a plain torch batched server (baseline/serving_bench.py)
        try:
            from baseline.serving_bench import run_serving_baseline

            base = run_serving_baseline(kind, dev, min(args.steps, 20), args.warmup, args.batch, ctx.rank)
            if ctx.world_size > 1:
                t = torch.tensor([base["value"]], dtype=torch.float64, device=dev)
                torch.distributed.all_reduce(t)
                base["value_all_ranks"] = float(t)
        except Exception as e:
            base = {"impl": "torch-baseline-server", "error": repr(e)[:300]}
        torch.cuda.empty_cache()
    if kind == "resnet50":
        import torchvision

        model = torch.nn.Sequential(torchvision.models.resnet50(num_classes=1000), torch.nn.Softmax(-1))
        conf = {"data": {"image_shape": "3,224,224", "filter": "topN(5)"}, "params": {"batch_size": bs, "dtype": os.environ.get("ZOO_SERVING_DTYPE", "bf16")}}
    else:
        from zoo.models.recommendation import NeuralCF

        model = NeuralCF(1_000_000, 100_000, 5, user_embed=64, item_embed=64, hidden_layers=[128, 64, 32], mf_embed=32, sparse=False)
        conf = {"data": {"filter": "topN(1)"}, "params": {"batch_size": bs, "dtype": os.environ.get("ZOO_SERVING_DTYPE", "fp32")}}
    # ZOO_SERVING_SHARED_BROKER=1: the reference's deployment shape -- ONE broker (hosted by rank 0, RESP socket), every rank's engine
    # a consumer of the same group, every rank's feeder a producer.  Default: one in-process broker per engine (replicas).
    shared = os.environ.get("ZOO_SERVING_SHARED_BROKER", "0") == "1" and ctx.world_size > 1
    if shared:
        from zoo.native.rt import Broker, RespClient

        port = [None]
        if ctx.rank == 0:
            host_broker = Broker(0)
            port[0] = host_broker.serve("127.0.0.1", 0)
        torch.distributed.broadcast_object_list(port, src=0)
        broker = host_broker if ctx.rank == 0 else RespClient("127.0.0.1", port[0])
        srv = ClusterServing(conf, model=model, device=dev, broker=broker, consumer=f"engine-{ctx.rank}").start()
        feed_db = host_broker if ctx.rank == 0 else RespClient("127.0.0.1", port[0])       # producers use their own connection
        feed_prog = host_broker if ctx.rank == 0 else RespClient("127.0.0.1", port[0])     # and so does the progress poll
    else:
        srv = ClusterServing(conf, model=model, device=dev).start()
        feed_db = srv.db
    q = InputQueue(conf, broker=feed_db)
    K, Wm = args.steps, max(args.warmup, 3)
    total = (K + Wm) * bs
    rng = np.random.RandomState(ctx.rank)
    if kind == "resnet50":
        import cv2

        jpgs = [q.base64_encode_image(cv2.imencode(".jpg", rng.randint(0, 256, (224, 224, 3), dtype=np.uint8))[1]) for _ in range(64)]

    def feed():
        if kind == "resnet50":
            for i in range(total):
                feed_db.xadd("image_stream", {"uri": f"r{ctx.rank}_{i}", "image": jpgs[i % 64]})
            return
        q.enqueue_tensor("r0", np.array([1, 1], dtype=np.int64))          # one record through the per-record path
        chunk = 1024                                                        # then bulk producers (one native call per 1024 records)
        for s0 in range(1, total, chunk):
            n = min(chunk, total - s0)
            ids = np.stack([rng.randint(1, 1_000_001, n), rng.randint(1, 100_001, n)], 1).astype(np.int64)
            while q.enqueue_tensors([f"r{s0 + k}" for k in range(n)], ids) == 0:
                pass

    t = threading.Thread(target=feed, daemon=True)
    t.start()

    def wait_global(target, limit_s):
        """Engines of a shared broker take whatever they get: progress is the SUM over ranks.  The ranks publish their counts in
        a hash of the broker itself (no CUDA or NCCL call from this thread: the engine thread may be capturing a CUDA graph)."""
        end = time.time() + limit_s
        while time.time() < end:
            feed_prog.hset("bench_progress", str(ctx.rank), str(srv.total_records))
            got = feed_prog.hgetall("bench_progress")
            if sum(int(v) for v in got.values()) >= target:
                return True
            time.sleep(0.02)
        return False

    if shared:
        wait_global(Wm * bs * ctx.world_size, 600)
    else:
        srv.wait_until(Wm * bs, 600)
    torch.cuda.synchronize()
    if ctx.world_size > 1:
        torch.distributed.barrier()
    n0, t0, l0 = srv.total_records, time.perf_counter(), len(srv.latencies)
    if shared:
        wait_global(total * ctx.world_size, 1200)
    else:
        srv.wait_until(total, 1200)
    torch.cuda.synchronize()
    t1, n1 = time.perf_counter(), srv.total_records
    lat = np.array([d for _, d in srv.latencies[l0:]] or [0.0]) * 1e3
    res = torch.tensor([n1 - n0, t1 - t0], dtype=torch.float64, device=dev)
    if ctx.world_size > 1:
        cnt = res[:1].clone()
        torch.distributed.all_reduce(cnt)
        tmax = res[1:].clone()
        torch.distributed.all_reduce(tmax, op=torch.distributed.ReduceOp.MAX)
        res = torch.cat([cnt, tmax])
    srv.close()
    if ctx.rank == 0:
        print(json.dumps({"metric": f"Cluster Serving records/sec ({kind})", "value": float(res[0] / res[1]), "unit": "records/s",
                          "n_gpus": ctx.world_size, "steps": K, "warmup": Wm, "ms_per_step": float(res[1]) * 1e3 / K,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": conf["params"]["dtype"], "data": "synthetic",
                          "impl": "zoo-b200", "p50_batch_ms": float(np.percentile(lat, 50)), "p99_batch_ms": float(np.percentile(lat, 99)),
                          "torch_baseline": base, "engine_stage_seconds": getattr(srv, "stage_s", None),
                          "jpeg_backend": getattr(getattr(srv, "_jpeg", None), "backend", None),
                          "vs_torch_baseline": (float(res[0] / res[1]) / (base.get("value_all_ranks") or base["value"]))
                          if base and base.get("value") else None,
                          "config": {"model": kind, "global_batch": bs * ctx.world_size, "seq_len": None,
                                     "parallelism": (f"{ctx.world_size} engines in one consumer group of ONE broker (rank 0 hosts it; RESP socket)" if shared
                                                     else f"{ctx.world_size} serving replicas (one engine + in-process broker per GPU)"),
                                     "timing": "wall clock over the whole pipeline: enqueue -> broker -> decode -> K13 -> model -> K14 -> result hash"}}))
    return 0
