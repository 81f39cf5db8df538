"""XShards: read csv files into shards, preprocess them in parallel on the worker pool, train from the shards."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import pandas as pd
import torch
import torch.nn as nn

import zoo.orca.data.pandas as zpd
from zoo.orca.learn.pytorch import Estimator
from zoo.ray import RayContext

if __name__ == "__main__":
    d = tempfile.mkdtemp()
    rng = np.random.RandomState(0)
    for i in range(4):
        x = rng.randn(500, 3)
        pd.DataFrame({"a": x[:, 0], "b": x[:, 1], "c": x[:, 2], "y": (x.sum(1) > 0).astype(float)}).to_csv(os.path.join(d, f"part-{i}.csv"), index=False)
    ctx = RayContext(num_workers=2).init()
    shards = zpd.read_csv(d, ctx)
    print("partitions:", shards.num_partitions(), "rows:", len(shards))

    def scale(df):
        df = df.copy()
        df[["a", "b", "c"]] = (df[["a", "b", "c"]] - df[["a", "b", "c"]].mean()) / df[["a", "b", "c"]].std()
        return df

    shards = shards.transform_shard(scale)
    model = nn.Sequential(nn.Linear(3, 8), nn.ReLU(), nn.Linear(8, 1), nn.Sigmoid())
    est = Estimator.from_torch(model=model, optimizer=torch.optim.Adam(model.parameters(), lr=1e-2), loss=nn.BCELoss(), metrics=["accuracy"])
    est.fit(shards, epochs=5, batch_size=100, feature_cols=["a", "b", "c"], label_cols=["y"])
    print(est.evaluate(shards, batch_size=500, feature_cols=["a", "b", "c"], label_cols=["y"]))
    ctx.stop()
