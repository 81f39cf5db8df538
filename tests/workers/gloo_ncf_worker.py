"""torchrun worker (CPU, gloo): BASELINE config 1 -- NeuralCF on synthetic MovieLens-1M-shaped ratings
through `Estimator.from_torch`, world_size=2.  Checks that ranks stay bit-identical and that the loss falls."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ["ZOO_FORCE_CPU"] = "1"
import numpy as np
import torch
import torch.distributed as dist

from zoo.common.nncontext import init_nncontext
from zoo.models.recommendation import NeuralCF
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.estimator import Estimator

ctx = init_nncontext()
assert ctx.world_size == 2 and ctx.backend == "gloo", ctx
torch.manual_seed(0)
U, I, C = 6040, 3706, 5
rng = np.random.RandomState(0)
n = 8192
x = np.stack([rng.randint(1, 301, n), rng.randint(1, 201, n)], 1).astype(np.int64)
y = ((x[:, 0] * 7 + x[:, 1] * 3) % C + 1).astype(np.int64)
model = NeuralCF(U, I, C, user_embed=20, item_embed=20, hidden_layers=[40, 20, 10], mf_embed=20)
est = Estimator.from_torch(model=model, optimizer=torch.optim.Adam(model.parameters(), lr=5e-3),
                           loss=SparseCategoricalCrossEntropy(zero_based_label=False), metrics=["accuracy"])
first = est.fit((x, y), epochs=1, batch_size=512)["loss"]
last = est.fit((x, y), epochs=5, batch_size=512)["loss"]
w = est.engine.fopt.arena.flat_p.clone()
w0 = w.clone()
dist.broadcast(w0, src=0)
assert torch.equal(w, w0), "ranks diverged"
acc = est.evaluate((x, y), batch_size=1024)["Top1Accuracy"]
assert last < first, (first, last)
pred = est.predict(x[:16])
assert pred.shape == (16, 5)
if ctx.rank == 0:
    print(f"GLOO_NCF_OK first={first:.4f} last={last:.4f} acc={acc:.3f}")
dist.barrier()
