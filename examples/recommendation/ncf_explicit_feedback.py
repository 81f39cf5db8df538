"""NeuralCF on MovieLens-1M-shaped explicit ratings (6,040 users x 3,706 items, ratings 1..5)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.common.nncontext import init_nncontext
from zoo.models.recommendation import NeuralCF, UserItemFeature
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.api.keras.optimizers import Adam

ap = argparse.ArgumentParser()
ap.add_argument("--epochs", type=int, default=2)
ap.add_argument("--batch-size", type=int, default=8000)
ap.add_argument("--ratings", type=int, default=200_000)
a = ap.parse_args()
init_nncontext("NCF example")
rng = np.random.RandomState(0)
users, items = rng.randint(1, 6041, a.ratings), rng.randint(1, 3707, a.ratings)
labels = ((users * 7 + items * 13) % 5 + 1).astype(np.int64)            # a learnable synthetic rating rule
x = np.stack([users, items], 1).astype(np.int64)
split = int(0.8 * a.ratings)
ncf = NeuralCF(user_count=6040, item_count=3706, class_num=5, hidden_layers=[20, 10], include_mf=False)
ncf.compile(optimizer=Adam(lr=1e-3), loss=SparseCategoricalCrossEntropy(zero_based_label=False), metrics=["accuracy"])
ncf.fit(x[:split], labels[:split], batch_size=a.batch_size, nb_epoch=a.epochs, validation_data=(x[split:], labels[split:]))
print("validation [loss, accuracy]:", ncf.evaluate(x[split:], labels[split:], batch_size=a.batch_size))
pairs = [UserItemFeature(int(u), int(i), x[k]) for k, (u, i) in enumerate(x[split:split + 2000])]
for r in ncf.recommend_for_user(pairs, 3)[:6]:
    print(r)
