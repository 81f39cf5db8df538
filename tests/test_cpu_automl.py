"""AutoML / Zouwu on CPU (reference tests: pyzoo/test/zoo/automl/**, pyzoo/test/zoo/zouwu/*.py): synthetic
DataFrames, smoke recipes, feature-transformer golden checks, save/restore round trips."""
import numpy as np
import pandas as pd
import pytest
import torch


@pytest.fixture(autouse=True)
def _cpu_ctx(monkeypatch):
    import zoo.common.nncontext as nc

    monkeypatch.setenv("ZOO_FORCE_CPU", "1")
    nc._CONTEXT = None
    yield
    nc._CONTEXT = None


def _df(n=200, extra=False, seed=0):
    rng = np.random.RandomState(seed)
    t = np.arange(n)
    df = pd.DataFrame({"datetime": pd.date_range("2019-01-01", periods=n, freq="h"),
                       "value": np.sin(t / 6.0) * 5 + 20 + rng.randn(n) * 0.1})
    if extra:
        df["extra"] = rng.rand(n)
    return df


def test_metrics():
    from zoo.automl.common.metrics import Evaluator

    y, p = np.array([[1.0, 2.0], [3.0, 4.0]]), np.array([[1.5, 2.0], [2.0, 5.0]])
    np.testing.assert_allclose(Evaluator.evaluate("mse", y, p), [(0.25 + 1) / 2, 0.5])
    assert abs(Evaluator.evaluate("mae", y, p, multioutput="uniform_average") - 0.625) < 1e-12
    np.testing.assert_allclose(Evaluator.evaluate("rmse", y, p), np.sqrt([0.625, 0.5]))
    assert Evaluator.evaluate("r2", y, y, "uniform_average") == 1.0
    assert Evaluator.evaluate("smape", y, y, "uniform_average") < 1e-6
    for m in ("me", "msle", "mpe", "mape", "mspe", "mdape", "smdape"):
        assert np.isfinite(Evaluator.evaluate(m, y, p)).all()
    assert Evaluator.get_metric_mode("r2") == "max" and Evaluator.get_metric_mode("mse") == "min"
    with pytest.raises(ValueError):
        Evaluator.evaluate("bogus", y, p)


def test_feature_transformer_golden(tmp_path):
    from zoo.automl.feature.time_sequence import TimeSequenceFeatureTransformer

    df = _df(50, extra=True)
    ft = TimeSequenceFeatureTransformer(future_seq_len=2, extra_features_col=["extra"])
    feats = ft.get_feature_list(df)
    assert feats == ["MONTH(datetime)", "WEEKDAY(datetime)", "DAY(datetime)", "HOUR(datetime)", "IS_WEEKEND(datetime)",
                     "IS_AWAKE(datetime)", "IS_BUSY_HOURS(datetime)", "extra"]
    x, y = ft.fit_transform(df, selected_features=["HOUR(datetime)", "IS_BUSY_HOURS(datetime)", "extra"], past_seq_len=4)
    assert x.shape == (50 - 4 - 2 + 1, 4, 4) and y.shape == (45, 2)
    v = (df["value"].values - df["value"].mean()) / df["value"].std(ddof=0)
    np.testing.assert_allclose(x[3, :, 0], v[3:7], rtol=1e-5)            # column 0 is the scaled target
    np.testing.assert_allclose(y[3], v[7:9], rtol=1e-5)
    h = df["datetime"].dt.hour.values.astype(float)
    np.testing.assert_allclose(x[0, :, 1], ((h - h.mean()) / h.std())[:4], rtol=1e-5)
    xt, _ = ft.transform(df, is_train=False)
    assert xt.shape == (47, 4, 4)
    out = ft.post_processing(df, np.zeros((47, 2)), is_train=False)
    assert list(out.columns) == ["datetime", "value_0", "value_1"] and len(out) == 47
    assert out["datetime"].iloc[0] == df["datetime"].iloc[4]               # first forecast follows the first window
    np.testing.assert_allclose(out["value_0"], df["value"].mean())         # scaled 0 == mean
    ft.save(str(tmp_path / "ft.json"), replace=True)
    ft2 = TimeSequenceFeatureTransformer.load(str(tmp_path / "ft.json"))
    x2, y2 = ft2.transform(df, is_train=True)
    np.testing.assert_allclose(x2, x)
    with pytest.raises(ValueError):
        ft.fit_transform(df.assign(datetime=df["datetime"].astype(str)), selected_features=[], past_seq_len=2)
    with pytest.raises(ValueError):
        ft.fit_transform(df.iloc[:3], selected_features=[], past_seq_len=4)
    with pytest.raises(Exception):
        TimeSequenceFeatureTransformer().transform(df)


def test_search_space_expansion():
    from zoo.automl.config.recipe import GridRandomRecipe, GridSearch, MTNetSmokeRecipe, RandomSample, BayesRecipe, convert_bayes_configs
    from zoo.automl.search.RayTuneSearchEngine import expand_search_space

    cfgs = expand_search_space({"a": GridSearch([1, 2]), "b": GridSearch(["x", "y"]), "c": 3,
                                "d": RandomSample(lambda spec: spec.config.a * 10)}, num_samples=2)
    assert len(cfgs) == 8 and all(c["d"] == c["a"] * 10 and c["c"] == 3 for c in cfgs)
    m = expand_search_space(MTNetSmokeRecipe().search_space(["f"]), 1)[0]
    assert m["past_seq_len"] == (m["long_num"] + 1) * m["time_step"]
    g = expand_search_space(GridRandomRecipe(look_back=(3, 5)).search_space(["a", "b", "c", "d"]), 1)
    assert len(g) == 16 and all(3 <= c["past_seq_len"] <= 5 for c in g)
    with pytest.raises(ValueError):
        GridRandomRecipe(look_back=1)
    b = BayesRecipe(look_back=(2, 4))
    sp = b.search_space(["f1", "f2"])
    assert sp["lstm_1_units_float"] == (8, 128) and "bayes_feature_f1" in sp and b.search_algorithm() == "BayesOpt"
    c = convert_bayes_configs({"bayes_feature_f1": 0.9, "bayes_feature_f2": 0.1, "lstm_1_units_float": 31.6, "dropout_1": 0.3})
    assert c == {"selected_features": ["f1"], "lstm_1_units": 32, "dropout_1": 0.3}


@pytest.mark.parametrize("model,extra", [("LSTM", {"lstm_1_units": 8, "lstm_2_units": 4}), ("Seq2seq", {"latent_dim": 8}),
                                         ("MTNet", {"time_step": 3, "long_num": 2, "ar_size": 2, "cnn_hid_size": 8, "rnn_hid_sizes": [6]}),
                                         ("TCN", {"hidden_units": 8, "levels": 2})])
def test_models_fit_save_restore(tmp_path, model, extra):
    from zoo.automl.model.time_sequence import TimeSequenceModel

    torch.manual_seed(0)
    np.random.seed(0)
    rng = np.random.RandomState(0)
    T = 9
    x = rng.randn(256, T, 3).astype(np.float32)
    fsl = 2 if model == "Seq2seq" else 1
    y = np.stack([x[:, -1, 0] * 0.5 + x[:, -2, 0] * 0.3] * fsl, 1).astype(np.float32)
    m = TimeSequenceModel(future_seq_len=fsl)
    cfg = dict(model=model, lr=0.01, batch_size=64, epochs=3, dropout=0.0, dropout_1=0.0, dropout_2=0.0, cnn_dropout=0.0, rnn_dropout=0.0, **extra)
    first = m.fit_eval(x, y if fsl > 1 else y[:, 0], **cfg)
    last = first
    for _ in range(6):
        last = m.fit_eval(x, y if fsl > 1 else y[:, 0], **cfg)
    assert last < first
    p = m.predict(x)
    assert p.shape == (256, fsl)
    mean, std = m.predict_with_uncertainty(x[:8], n_iter=3)
    assert mean.shape == (8, fsl) and (std >= 0).all()
    m.save(str(tmp_path / "m.h5"), str(tmp_path / "c.json"))
    from zoo.automl.common.util import load_config

    m2 = TimeSequenceModel().restore(str(tmp_path / "m.h5"), **load_config(str(tmp_path / "c.json")))
    np.testing.assert_allclose(m2.predict(x), p, atol=1e-6)


def test_reference_state_dict_layout():
    """`LSTMModel` keeps the reference's parameter names so its torch checkpoints load as they are."""
    from zoo.automl.model.VanillaLSTM_pytorch import LSTMModel

    keys = set(LSTMModel(3, 4, 2, 0.1, 1).state_dict())
    assert keys == {f"lstm.{w}_l{l}" for w in ("weight_ih", "weight_hh", "bias_ih", "bias_hh") for l in (0, 1)} | {"fc.weight", "fc.bias"}


def test_autots_trainer_end_to_end(tmp_path):
    from zoo.automl.config.recipe import LSTMGridRandomRecipe
    from zoo.zouwu.autots import AutoTSTrainer, TSPipeline

    train, val = _df(300), _df(120, seed=1)
    trainer = AutoTSTrainer(horizon=1, dt_col="datetime", target_col="value")
    recipe = LSTMGridRandomRecipe(num_rand_samples=1, epochs=2, training_iteration=2, look_back=6, lstm_1_units=[8], lstm_2_units=[4, 8], batch_size=[32])
    ppl = trainer.fit(train, val, metric="mse", recipe=recipe)
    assert isinstance(ppl, TSPipeline)
    mse, smape = ppl.evaluate(val, metrics=["mse", "smape"])
    assert mse[0] < 30 and smape[0] < 20
    pred = ppl.predict(val)
    assert list(pred.columns) == ["datetime", "value"] and len(pred) == 120 - 6 + 1
    f = ppl.save(str(tmp_path / "ts.ppl"))
    ppl2 = TSPipeline.load(f)
    np.testing.assert_allclose(ppl2.predict(val)["value"].values, pred["value"].values, rtol=1e-5)
    before = ppl2.evaluate(val, metrics=["mse"])[0][0]
    ppl2.fit(train, val, epochs=3)                               # incremental fit
    assert ppl2.evaluate(val, metrics=["mse"])[0][0] < before * 1.5
    out, sigma = ppl2.predict_with_uncertainty(val, n_iter=3)
    assert len(out) == len(sigma) == len(pred)
    ppl2.fit(train, val, epochs=2, model="TCN", past_seq_len=8, hidden_units=8, levels=2, lr=0.01, batch_size=32)   # fixed-config re-train
    assert len(ppl2.predict(val)) == 120 - 8 + 1


def test_parallel_search_on_worker_pool():
    from zoo.automl.config.recipe import LSTMGridRandomRecipe
    from zoo.automl.regression.time_sequence_predictor import TimeSequencePredictor
    from zoo.ray import RayContext

    ctx = RayContext(num_workers=2).init()
    try:
        tsp = TimeSequencePredictor(future_seq_len=2, logs_dir="/tmp/zoo_automl_test")
        recipe = LSTMGridRandomRecipe(epochs=1, training_iteration=1, look_back=4, lstm_1_units=[4, 8], lstm_2_units=[4], batch_size=[64])
        ppl = tsp.fit(_df(150), _df(80, seed=2), metric="mae", recipe=recipe)
        assert ppl.config["future_seq_len"] == 2 and ppl.config["lstm_1_units"] in (4, 8)
        assert tsp.predict(_df(40, seed=3)).shape == (40 - 4 + 1, 3)
    finally:
        ctx.stop()


def test_forecasters():
    from zoo.zouwu.model import LSTMForecaster, MTNetForecaster, TCNForecaster

    torch.manual_seed(0)
    rng = np.random.RandomState(0)
    x = rng.randn(400, 6, 2).astype(np.float32)
    y = (x[:, -1, 0] * 0.7 - x[:, -2, 1] * 0.2).reshape(-1, 1)
    for f in (LSTMForecaster(target_dim=1, feature_dim=2, lstm_1_units=8, lstm_2_units=4, dropout_1=0.0, dropout_2=0.0, lr=0.01),
              TCNForecaster(target_dim=1, feature_dim=2, num_channels=(8, 8), dropout=0.0, lr=0.01),
              MTNetForecaster(target_dim=1, feature_dim=2, long_series_num=2, series_length=2, ar_window_size=1, cnn_height=1, cnn_hid_size=8,
                              rnn_hid_sizes=(6,), cnn_dropout=0.0, rnn_dropout=0.0, lr=0.01)):
        before = f.evaluate(x, y, batch_size=200)[0]
        f.fit(x, y, batch_size=50, epochs=6)
        after = f.evaluate(x, y, batch_size=200)[0]
        assert after < before, type(f).__name__
        assert f.predict(x[:7]).shape == (7, 1)
        m, s = f.predict_with_uncertainty(x[:5], n_iter=3)
        assert m.shape == s.shape == (5, 1)
    mt = MTNetForecaster(feature_dim=2, long_series_num=2, series_length=2)
    lt, st = mt.preprocess_input(x)
    assert lt.shape == (400, 2, 2, 2) and st.shape == (400, 2, 2)
    with pytest.raises(ValueError):
        mt.preprocess_input(x[:, :5])


def test_keras_recipe_forecasters():
    """`VanillaLSTM` / `LSTMSeq2Seq` with the reference's Keras-recipe architecture, defaults and training scheme
    (pyzoo/zoo/automl/model/VanillaLSTM.py:50-90, Seq2Seq.py:50-165), save / restore round trip."""
    import os
    import tempfile

    from zoo.automl.model.Seq2Seq import LSTMSeq2Seq
    from zoo.automl.model.time_sequence import TimeSequenceModel
    from zoo.automl.model.VanillaLSTM import VanillaLSTM

    torch.manual_seed(0)
    rng = np.random.RandomState(0)
    t = np.arange(400, dtype=np.float32)
    series = np.sin(t / 6.0) + 0.05 * rng.randn(400).astype(np.float32)
    past, fut = 12, 3
    x = np.stack([np.stack([series[i:i + past], np.cos(t[i:i + past] / 6.0)], -1) for i in range(380)]).astype(np.float32)
    y1 = np.stack([series[i + past] for i in range(380)]).astype(np.float32)
    y3 = np.stack([series[i + past:i + past + fut] for i in range(380)]).astype(np.float32)

    m = VanillaLSTM(future_seq_len=1)
    first = m.fit_eval(x, y1, epochs=1, batch_size=64, lr=0.01)
    last = m.fit_eval(x, y1, epochs=15, batch_size=64, lr=0.01)
    assert last < first and last < 0.1, (first, last)
    assert type(m.optimizer).__name__ == "RMSprop" and m.model.lstm1.hidden_size == 20 and m.model.lstm2.hidden_size == 10
    mean, std = m.predict_with_uncertainty(x[:16], n_iter=5)
    assert mean.shape == (16, 1) and (std > 0).any()

    s = LSTMSeq2Seq(future_seq_len=fut)
    first = s.fit_eval(x, y3, epochs=1, batch_size=64, lr=0.01, latent_dim=32, dropout=0.0)
    last = s.fit_eval(x, y3, epochs=15, batch_size=64, lr=0.01, latent_dim=32, dropout=0.0)
    assert last < first and last < 0.15, (first, last)
    p = s.predict(x[:10])
    assert p.shape == (10, fut)
    # teacher-forced training output and step-wise decoding agree when the fed-back values are the truth
    with torch.no_grad():
        xb, yb = torch.from_numpy(x[:4]), torch.from_numpy(s._prepare_y(y3[:4]))
        s.model.eval()
        tf = s.model(xb, torch.cat([xb[:, -1:, :1], yb[:, :-1]], 1))
        assert torch.allclose(tf[:, 0], s.model(xb)[:, 0], atol=1e-5)          # the first step sees the same input either way
    with tempfile.TemporaryDirectory() as d:
        s.save(os.path.join(d, "m.pt"), os.path.join(d, "c.json"))
        from zoo.automl.common.util import load_config

        r = LSTMSeq2Seq(future_seq_len=fut).restore(os.path.join(d, "m.pt"), **load_config(os.path.join(d, "c.json")))
        assert np.allclose(r.predict(x[:10]), p, atol=1e-6)

    # two target columns: y [n, future, 2]
    y2c = np.stack([y3, -y3], -1)
    x2c = np.concatenate([x[..., :1], -x[..., :1], x[..., 1:]], -1)
    s2 = LSTMSeq2Seq(future_seq_len=fut)
    s2.fit_eval(x2c, y2c, epochs=2, batch_size=64, latent_dim=16)
    assert s2.predict(x2c[:5]).shape == (5, fut, 2) and s2.model.decoder_dense.out_features == 2

    ts = TimeSequenceModel(future_seq_len=fut)
    ts.fit_eval(x, y3, model="Seq2seq_keras", epochs=1, latent_dim=8)
    assert isinstance(ts.model, LSTMSeq2Seq) and ts.predict(x[:3]).shape == (3, fut)
