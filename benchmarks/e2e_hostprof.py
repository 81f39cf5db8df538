"""Host-side profile (cProfile) of the end-to-end fit loop on the NCF bench config: where does the Python time per step go?
usage: python benchmarks/e2e_hostprof.py [steps]"""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import build_ncf
from zoo.common.nncontext import init_nncontext
from zoo.feature.common import FeatureSet
from zoo.pipeline.estimator import Estimator


class A:
    config, scale, batch = "ncf", 0.2, 0


steps = int(sys.argv[1]) if len(sys.argv) > 1 else 400
ctx = init_nncontext()
cfg = build_ncf(A, ctx, torch)
opt = torch.optim.Adam([p for p in cfg["model"].parameters() if p.requires_grad] or [torch.zeros(1, requires_grad=True)], lr=1e-3)
est = Estimator.from_torch(model=cfg["model"], optimizer=opt, loss=cfg["loss"], ctx=ctx)
host = cfg["make_batches"](48, torch.device("cpu"))
X, Y = torch.cat([b[0] for b in host]), torch.cat([b[1] for b in host])
fs = FeatureSet(features=X, labels=Y, memory_type="DRAM", shuffle=False)
est.engine.async_loss = True
est.fit(fs, batch_size=cfg["batch"], max_steps=20, log_every=1, shard=False)      # warm-up + graph capture
torch.cuda.synchronize()
pr = cProfile.Profile()
t0 = time.time()
pr.enable()
est.fit(fs, batch_size=cfg["batch"], max_steps=20 + steps, log_every=1, shard=False)
torch.cuda.synchronize()
pr.disable()
dt = time.time() - t0
print(f"{steps} steps in {dt * 1e3:.1f} ms -> {dt / steps * 1e6:.1f} us/step under cProfile")
pstats.Stats(pr).sort_stats("tottime").print_stats(28)
