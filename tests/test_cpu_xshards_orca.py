"""XShards, read_csv/json, RayContext worker pool, spawning Estimator (reference tests:
pyzoo/test/zoo/xshard/test_ray_pandas.py:38-91, pyzoo/test/zoo/ray/test_ray_on_local.py)."""
import json
import os

import numpy as np
import pandas as pd
import pytest
import torch
import torch.nn as nn


@pytest.fixture(autouse=True)
def _cpu_ctx(monkeypatch):
    import zoo.common.nncontext as nc

    monkeypatch.setenv("ZOO_FORCE_CPU", "1")
    nc._CONTEXT = None
    yield
    nc._CONTEXT = None


def _write_csvs(d, n_files=4, rows=25):
    rng = np.random.RandomState(0)
    for i in range(n_files):
        pd.DataFrame({"user": rng.randint(1, 50, rows), "item": rng.randint(1, 30, rows), "label": rng.randint(1, 6, rows),
                      "value": rng.rand(rows)}).to_csv(os.path.join(d, f"part{i}.csv"), index=False)


def test_read_csv_apply_collect_repartition(tmp_path):
    from zoo.ray import RayContext
    from zoo.xshard.pandas import read_csv, read_json

    _write_csvs(str(tmp_path))
    ctx = RayContext(num_workers=2, local_mode=True).init()
    shards = read_csv(str(tmp_path), ctx)
    data = shards.collect()
    assert len(data) == 2 and sum(len(d) for d in data) == 100 and list(data[0].columns) == ["user", "item", "label", "value"]

    def negative(df, column_name):
        df[column_name] = df[column_name] * (-1)
        return df

    shards.apply(negative, "value")
    assert all((d["value"] <= 0).all() for d in shards.collect())
    shards.repartition(1)
    assert len(shards.get_partitions()) == 1 and len(shards.get_partitions()[0].get_data()) == 2
    one = read_csv(os.path.join(str(tmp_path), "part0.csv") + "," + os.path.join(str(tmp_path), "part1.csv"), ctx)
    assert sum(len(d) for d in one.collect()) == 50
    jp = tmp_path / "a.json"
    jp.write_text(json.dumps({"a": [1, 2, 3], "b": [4, 5, 6]}))
    assert len(read_json(str(jp), ctx).collect()[0]) == 3
    with pytest.raises(Exception):
        read_csv(str(tmp_path / "missing"), ctx)
    ctx.stop()


def test_xshards_partition_to_estimator():
    from zoo.orca.data import XShards
    from zoo.orca.learn.pytorch import Estimator

    rng = np.random.RandomState(0)
    x = rng.randn(1000, 4).astype(np.float32)
    y = (x.sum(1, keepdims=True) > 0).astype(np.float32)
    sh = XShards.partition({"x": x, "y": y}, num_shards=4)
    assert sh.num_partitions() == 4 and len(sh) == 1000
    sh2 = sh.transform_shard(lambda s: {"x": s["x"] * 2, "y": s["y"]})
    assert np.allclose(sh2.collect()[0]["x"], sh.collect()[0]["x"] * 2) and sh2 is not sh
    assert [len(s) for s in (sh.collect_local(0, 2), sh.collect_local(1, 2))] == [2, 2]
    model = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 1), nn.Sigmoid())
    est = Estimator.from_torch(model=model, optimizer=torch.optim.Adam(model.parameters(), lr=1e-2), loss=nn.BCELoss(), metrics=["accuracy"])
    est.fit(sh, epochs=5, batch_size=100)
    assert est.evaluate(sh, batch_size=200)["Top1Accuracy"] > 0.9
    assert est.predict(sh).shape == (1000, 1)
    df_sh = XShards([pd.DataFrame({"a": x[:500, 0], "b": x[:500, 1], "y": y[:500, 0]}), pd.DataFrame({"a": x[500:, 0], "b": x[500:, 1], "y": y[500:, 0]})])
    ds = df_sh.to_torch_dataset(feature_cols=["a", "b"], label_cols=["y"])
    assert len(ds) == 1000 and ds[0][0].shape == (2,)
    assert df_sh.split(5).num_partitions() == 5


class Counter:
    def __init__(self, start):
        self.v = start

    def add(self, k):
        self.v += k
        return self.v


def test_raycontext_pool_and_actor():
    from zoo.ray import RayContext

    ctx = RayContext(num_workers=2).init()
    assert RayContext.get() is ctx
    k = 7
    assert sorted(ctx.map(lambda a: a * k, [1, 2, 3])) == [7, 14, 21]          # closures are cloudpickled
    pids = set(ctx.map(lambda _: os.getpid(), range(8)))
    assert os.getpid() not in pids and 1 <= len(pids) <= 2
    a = ctx.actor(Counter, 10)
    assert [a.add.remote(1).get(), a.add.remote(5).get()] == [11, 16]          # state lives in the actor process
    a.kill()
    with pytest.raises(ZeroDivisionError):
        ctx.remote(lambda: 1 / 0).get()
    ctx.stop()
    with pytest.raises(Exception):
        RayContext.get()


def test_spawning_estimator_two_workers():
    """MXNetTrainer-shaped use: creators shipped to 2 workers, gloo data-parallel training."""
    from zoo.orca.learn.pytorch import Estimator

    def data_creator(config, batch_size):
        rng = np.random.RandomState(0)
        x = rng.randn(2048, 4).astype(np.float32)
        return x, (x.sum(1, keepdims=True) > 0).astype(np.float32)

    est = Estimator.from_torch(model_creator=lambda c: nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 1), nn.Sigmoid()),
                               optimizer_creator=lambda m, c: torch.optim.Adam(m.parameters(), lr=c["lr"]),
                               loss_creator=lambda c: nn.BCELoss(), metrics=["accuracy"], config={"lr": 1e-2}, workers_per_node=2)
    stats = est.fit(data_creator, epochs=4, batch_size=128)
    assert len(stats) == 2 and stats[0]["iterations"] == stats[1]["iterations"] == 64
    x, y = data_creator({}, 0)
    with torch.no_grad():
        acc = ((est.get_model()(torch.from_numpy(x)) > 0.5).float().numpy() == y).mean()
    assert acc > 0.9
    assert est.evaluate(data_creator, batch_size=256)["Top1Accuracy"] > 0.9


def test_nest():
    from zoo.util import nest

    s = {"b": [1, (2, 3)], "a": 0}
    assert nest.flatten(s) == [0, 1, 2, 3]
    assert nest.pack_sequence_as(s, [10, 11, 12, 13]) == {"a": 10, "b": [11, (12, 13)]}
    with pytest.raises(ValueError):
        nest.pack_sequence_as(s, [1, 2])


def test_remote_filesystem_scheme(tmp_path):
    """hdfs:// / s3:// handling goes through pyarrow file systems; a directory mounted as `mock://` stands in for a remote store."""
    import pandas as pd
    import pyarrow.fs as pafs

    from zoo.common.utils import get_file_list, get_remote_file_to_local, load_from_file, put_local_file_to_remote, save_file
    from zoo.util import filesystem as zfs
    from zoo.xshard.pandas import read_csv

    store = tmp_path / "store"
    store.mkdir()
    zfs.register_filesystem("mock", pafs.SubTreeFileSystem(str(store), pafs.LocalFileSystem()))
    try:
        for i in range(3):
            pd.DataFrame({"a": range(i * 10, i * 10 + 10), "b": 1.5}).to_csv(tmp_path / f"part{i}.csv", index=False)
            put_local_file_to_remote(str(tmp_path / f"part{i}.csv"), f"mock://data/part{i}.csv")
        assert get_file_list("mock://data") == [f"mock://data/part{i}.csv" for i in range(3)]
        assert zfs.exists("mock://data/part1.csv") and not zfs.exists("mock://data/nope.csv") and zfs.isdir("mock://data")
        shards = read_csv("mock://data")
        df = pd.concat(shards.collect(), ignore_index=True)
        assert len(df) == 30 and sorted(df["a"]) == list(range(30))
        get_remote_file_to_local("mock://data/part2.csv", str(tmp_path / "back.csv"))
        assert pd.read_csv(tmp_path / "back.csv")["a"].tolist() == list(range(20, 30))
        import pytest as _pt
        with _pt.raises(FileExistsError):
            put_local_file_to_remote(str(tmp_path / "part0.csv"), "mock://data/part0.csv")
        save_file(lambda p, obj: open(p, "w").write(obj), "mock://models/m.txt", obj="hello")
        assert load_from_file(lambda p: open(p).read(), "mock://models/m.txt") == "hello"
        with _pt.raises(OSError):
            get_file_list("hdfs://no-such-namenode:1/x")       # no Hadoop client on this box: a clear error, not a hang
    finally:
        zfs.unregister_filesystem("mock")
