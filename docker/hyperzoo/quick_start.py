"""Smoke job for a hyperzoo pod (reference: docker/hyperzoo/quick_start.py trains a small model through Spark on K8s).
Run as `submit.sh docker/hyperzoo/quick_start.py`: every visible GPU trains NCF for a few steps and rank 0 prints the throughput."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np

from zoo.common.nncontext import init_nncontext
from zoo.models.recommendation import NeuralCF

ctx = init_nncontext("hyperzoo quick start")
rng = np.random.RandomState(ctx.rank)
users, items, n = 20000, 5000, 200000
x = np.stack([rng.randint(1, users + 1, n), rng.randint(1, items + 1, n)], 1).astype(np.int64)
y = ((x[:, 0] + x[:, 1]) % 5).astype(np.int64)
model = NeuralCF(users, items, 5, user_embed=32, item_embed=32, hidden_layers=[64, 32, 16], mf_embed=16)
model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
model.fit(x, y, batch_size=4096 * ctx.world_size, nb_epoch=int(os.environ.get("EPOCHS", 2)))
if ctx.rank == 0:
    print("quick start done on %d rank(s); last throughput %.0f records/s" % (ctx.world_size, model._engine.last_throughput if model._engine else 0.0))
