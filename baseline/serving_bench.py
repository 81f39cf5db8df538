"""Same-box baseline for the Cluster Serving configs: a plain PyTorch batched inference server written the ordinary way -- a
Python queue of records, cv2 decode on a thread pool, torch ops for resize / normalise, the model under bf16 autocast, torch.topk,
results in a dict -- with the same feeder, batch size, warm-up and counting rules as the product arm (benchmarks/configs.py::_serving).
None of this repository's broker, kernels or engine is used.  Reference harness shape:
zoo/src/test/resources/serving/offline-benchmark (enqueue N images, read records/s from the log)."""
from __future__ import annotations

import queue
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch
import torch.nn.functional as F


def run_serving_baseline(kind, dev, steps, warmup, batch=0, rank=0):
    bs = batch or (64 if kind == "resnet50" else 4096)
    K, Wm = steps, max(warmup, 3)
    total = (K + Wm) * bs
    rng = np.random.RandomState(rank)
    q: queue.Queue = queue.Queue(maxsize=8 * bs)
    results, lat = {}, []
    if kind == "resnet50":
        import cv2
        import torchvision

        model = torch.nn.Sequential(torchvision.models.resnet50(num_classes=1000), torch.nn.Softmax(-1)).to(dev).eval()
        import base64

        # records arrive base64-encoded, as in the reference's client protocol (pyzoo/zoo/serving/client.py:87-126)
        jpgs = [base64.b64encode(cv2.imencode(".jpg", rng.randint(0, 256, (224, 224, 3), dtype=np.uint8))[1].tobytes()) for _ in range(64)]
        mean = torch.tensor([123.68, 116.779, 103.939], device=dev).view(1, 3, 1, 1)
        std = torch.tensor([58.393, 57.12, 57.375], device=dev).view(1, 3, 1, 1)
        pool = ThreadPoolExecutor(16)

        def feed():
            for i in range(total):
                q.put((f"r{i}", jpgs[i % 64], time.perf_counter()))

        def infer(items):
            imgs = list(pool.map(lambda b: cv2.imdecode(np.frombuffer(base64.b64decode(b), np.uint8), cv2.IMREAD_COLOR), [it[1] for it in items]))
            x = torch.from_numpy(np.stack(imgs)).to(dev, non_blocking=True).permute(0, 3, 1, 2).float()
            x = (F.interpolate(x, size=(224, 224), mode="bilinear", align_corners=False).flip(1) - mean) / std
            with torch.autocast("cuda", dtype=torch.bfloat16):
                p = model(x)
            v, i = torch.topk(p.float(), 5, dim=1)
            return v.cpu().numpy(), i.cpu().numpy()
    else:
        class NCF(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.mu, self.mi = torch.nn.Embedding(1_000_001, 64), torch.nn.Embedding(100_001, 64)
                self.gu, self.gi = torch.nn.Embedding(1_000_001, 32), torch.nn.Embedding(100_001, 32)
                self.mlp = torch.nn.Sequential(torch.nn.Linear(128, 128), torch.nn.ReLU(), torch.nn.Linear(128, 64), torch.nn.ReLU(),
                        torch.nn.Linear(64, 32), torch.nn.ReLU())
                self.last = torch.nn.Linear(64, 5)

            def forward(self, x):
                u, i = x[:, 0], x[:, 1]
                h = self.mlp(torch.cat([self.mu(u), self.mi(i)], 1))
                return torch.softmax(self.last(torch.cat([h, self.gu(u) * self.gi(i)], 1)), -1)

        model = NCF().to(dev).eval()

        def feed():
            for s0 in range(0, total, 1024):
                n = min(1024, total - s0)
                ids = np.stack([rng.randint(1, 1_000_001, n), rng.randint(1, 100_001, n)], 1).astype(np.int64)
                t = time.perf_counter()
                for k in range(n):
                    q.put((f"r{s0 + k}", ids[k], t))

        def infer(items):
            x = torch.from_numpy(np.stack([it[1] for it in items])).to(dev, non_blocking=True)
            p = model(x)
            v, i = torch.topk(p, 1, dim=1)
            return v.cpu().numpy(), i.cpu().numpy()

    threading.Thread(target=feed, daemon=True).start()
    done, t0, n0 = 0, None, 0
    with torch.no_grad():
        while done < total:
            items = [q.get()]
            while len(items) < bs:
                try:
                    items.append(q.get_nowait())
                except queue.Empty:
                    break                       # continuous batching: never wait for a full batch
            v, i = infer(items)
            now = time.perf_counter()
            for k, it in enumerate(items):
                results[it[0]] = (i[k].tolist(), v[k].tolist())
            done += len(items)
            if t0 is None and done >= Wm * bs:
                torch.cuda.synchronize()
                t0, n0 = time.perf_counter(), done
            elif t0 is not None:
                lat.append((now - items[0][2]) * 1e3)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    lat = np.array(lat) if lat else np.zeros(1)
    return {"impl": "torch-baseline-server", "value": (done - n0) / dt, "unit": "records/s", "records": done - n0, "seconds": dt,
            "p50_batch_ms": float(np.percentile(lat, 50)), "p99_batch_ms": float(np.percentile(lat, 99)), "batch_size": bs}
