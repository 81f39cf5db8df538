#!/usr/bin/env python
"""Synthetic stand-ins for the datasets the reference downloads (scripts/data/{movielens-1m,news20,glove,NAB}/get_*.sh):
there is no network on the box.  usage: gen_synthetic.py movielens|census|nab|glove|news20|hics OUT_DIR"""
import os
import sys

import numpy as np
import pandas as pd

kind, out = sys.argv[1], sys.argv[2]
os.makedirs(out, exist_ok=True)
rng = np.random.RandomState(0)
if kind == "movielens":
    n = 1_000_209
    u, i = rng.randint(1, 6041, n), rng.randint(1, 3707, n)
    pd.DataFrame({"user": u, "item": i, "rating": (u * 7 + i * 13) % 5 + 1, "ts": rng.randint(956703932, 1046454590, n)}).to_csv(
        os.path.join(out, "ratings.dat"), sep=":", header=False, index=False)
elif kind == "census":
    n = 32561
    pd.DataFrame({"age": rng.randint(17, 90, n), "workclass": rng.randint(0, 9, n), "education": rng.randint(0, 16, n), "education_num": rng.randint(1, 17, n),
                  "marital_status": rng.randint(0, 7, n), "occupation": rng.randint(0, 15, n), "relationship": rng.randint(0, 6, n),
                  "capital_gain": rng.exponential(1000, n), "capital_loss": rng.exponential(80, n), "hours_per_week": rng.randint(1, 99, n),
                  "income_bracket": rng.randint(0, 2, n)}).to_csv(os.path.join(out, "adult.data"), index=False)
elif kind == "nab":
    t = np.arange(10320)
    pd.DataFrame({"timestamp": pd.date_range("2014-07-01", periods=len(t), freq="30min"),
                  "value": 15000 + 8000 * np.sin(2 * np.pi * t / 48) + 800 * rng.randn(len(t))}).to_csv(os.path.join(out, "nyc_taxi.csv"), index=False)
elif kind == "glove":
    with open(os.path.join(out, "glove.6B.50d.txt"), "w") as f:
        for k in range(20000):
            f.write(f"word{k} " + " ".join(f"{v:.5f}" for v in rng.randn(50) * 0.4) + "\n")
elif kind == "news20":
    cats = ["alt.atheism", "comp.graphics", "rec.autos", "sci.space", "talk.politics.misc"]
    words = [f"w{k}" for k in range(400)]
    for ci, c in enumerate(cats):
        os.makedirs(os.path.join(out, c), exist_ok=True)
        for d in range(60):                                  # a post: words drawn mostly from the category's own slice of the vocabulary
            own = rng.choice(words[ci * 60:(ci + 1) * 60], 50)
            shared = rng.choice(words[300:], 30)
            with open(os.path.join(out, c, str(10000 + d)), "w") as f:
                f.write(" ".join(rng.permutation(np.concatenate([own, shared]))))
elif kind == "hics":
    n, dims = 1000, 100
    x = rng.rand(n, dims)
    for s0 in range(0, 20, 4):                               # five 4-dimensional subspaces with clustered (correlated) values
        centre = rng.rand(3, 4)
        x[:, s0:s0 + 4] = centre[rng.randint(0, 3, n)] + 0.03 * rng.randn(n, 4)
    outlier = np.zeros(n, dtype=int)
    idx = rng.choice(n, 20, replace=False)
    for i in idx:                                            # an outlier breaks the correlation inside one subspace only
        s0 = 4 * rng.randint(0, 5)
        x[i, s0:s0 + 4] = rng.rand(4)
        outlier[i] = 1
    cols = [f"var_{k:04d}" for k in range(dims)]
    pd.DataFrame(np.column_stack([x, outlier]), columns=cols + ["class"]).to_csv(os.path.join(out, "synth_multidim_100_000.csv"), sep=";", index=False)
else:
    sys.exit("unknown dataset " + kind)
print("wrote", out)
