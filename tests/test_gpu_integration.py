"""Higher layers on the GPU kernels: keras-style models (tcgen05 Dense, K2 embeddings, K6 recurrent) through fit,
Cluster Serving end to end (K13 / bf16 model / K14), InferenceModel CUDA-graph replay, AutoML models (K6 / K7) and
NNFrames.  Numerics are checked against the same models run on CPU."""
import copy

import numpy as np
import pandas as pd
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _ctx():
    import zoo.common.nncontext as nc

    nc._CONTEXT = None
    yield
    nc._CONTEXT = None


def test_keras_sequential_on_gpu_matches_cpu_and_trains():
    from zoo.native import lib as nl
    from zoo.pipeline.api.keras.engine.topology import Sequential
    from zoo.pipeline.api.keras.layers import GRU, LSTM, Dense, Dropout, Embedding

    torch.manual_seed(0)
    m = Sequential([Embedding(50, 16, input_shape=(12,)), LSTM(24, return_sequences=True), GRU(16), Dense(32, activation="relu"), Dropout(0.0),
                    Dense(3, activation="softmax")])
    x = torch.randint(0, 50, (64, 12))
    cpu = m(x).detach()
    g = copy.deepcopy(m).cuda()
    n0 = nl.LAUNCH_COUNT["n"]
    gpu = g(x.cuda()).detach().float().cpu()
    assert nl.LAUNCH_COUNT["n"] - n0 >= 5          # embedding gather, 2 recurrent kernels, 2 dense GEMMs
    torch.testing.assert_close(gpu, cpu, atol=2e-2, rtol=5e-2)
    rng = np.random.RandomState(0)
    xs = rng.randint(0, 50, (2000, 12))
    ys = (xs[:, -1] % 3).astype(np.int64)
    from zoo.pipeline.api.keras.optimizers import Adam

    m2 = Sequential([Embedding(50, 16, input_shape=(12,)), LSTM(24), Dense(3, activation="softmax")])
    m2.compile(optimizer=Adam(lr=1e-2), loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    m2.fit(xs, ys, batch_size=200, nb_epoch=12)
    assert next(m2.parameters()).is_cuda
    assert m2.evaluate(xs, ys, batch_size=500)[1] > 0.9


def test_serving_on_gpu_matches_direct_inference():
    from zoo.ops.serving_ops import preprocess_reference
    from zoo.serving import ClusterServing, InputQueue, OutputQueue

    torch.manual_seed(0)
    model = nn.Sequential(nn.Conv2d(3, 8, 3, stride=4), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(8, 20), nn.Softmax(-1))
    conf = {"data": {"image_shape": "3,64,64", "filter": "topN(3)"}, "params": {"batch_size": 16, "dtype": "fp32"}}
    srv = ClusterServing(conf, model=copy.deepcopy(model)).start()
    assert srv.device.type == "cuda"
    inq, outq = InputQueue(conf, broker=srv.db), OutputQueue(conf, broker=srv.db)
    rng = np.random.RandomState(0)
    imgs = {f"i{k}": rng.randint(0, 256, (64, 64, 3), dtype=np.uint8) for k in range(40)}
    import base64

    import cv2

    for uri, im in imgs.items():
        inq.enqueue_image(uri, im)
    assert srv.wait_until(40, 60)
    res = outq.dequeue()
    srv.close()
    assert set(res) == set(imgs)
    # direct path: JPEG round trip on the host, reference preprocessing, the same model on CPU
    for uri in ("i0", "i17", "i39"):
        dec = cv2.imdecode(np.frombuffer(cv2.imencode(".jpg", imgs[uri])[1], np.uint8), cv2.IMREAD_COLOR)
        x = preprocess_reference([dec], (64, 64), None, srv.conf.mean, srv.conf.std, swap_rb=True)
        p = model(x)[0]
        top = p.topk(3)
        got = eval(res[uri])
        # the engine decodes with nvJPEG on the GPU, this check with OpenCV on the host: the two decoders differ by a grey level or
        # two (IDCT rounding, chroma upsampling), which moves the probabilities in the third decimal -- classes whose probabilities
        # are closer than that may swap places, so every returned (class, probability) is checked against the direct model output
        assert got[0][0] == int(top.indices[0]) or abs(float(p[got[0][0]]) - float(top.values[0])) < 5e-3
        for cls, prob in got:
            assert abs(prob - float(p[cls])) < 5e-3, (uri, got, top)
        assert min(g[1] for g in got) >= float(top.values[2]) - 5e-3


def test_inference_model_graph_replay_matches_eager():
    from zoo.pipeline.inference import InferenceModel

    torch.manual_seed(0)
    net = nn.Sequential(nn.Linear(32, 64), nn.ReLU(), nn.Linear(64, 8))
    im = InferenceModel(supported_concurrent_num=1).load_torch(copy.deepcopy(net))
    xs = [np.random.RandomState(i).rand(16, 32).astype(np.float32) for i in range(6)]
    outs = [im.predict(x) for x in xs]                   # from the 3rd call on this shape is a CUDA-graph replay
    slot = im._slots.get()
    assert slot.graphs, "recurring shape should have been captured"
    im._slots.put(slot)
    for x, o in zip(xs, outs):
        np.testing.assert_allclose(o, net(torch.from_numpy(x)).detach().numpy(), atol=1e-4, rtol=1e-4)
    assert im.predict(np.random.rand(5, 32).astype(np.float32)).shape == (5, 8)     # a new shape still works


@pytest.mark.parametrize("model,extra", [("LSTM", {"lstm_1_units": 16, "lstm_2_units": 8}), ("Seq2seq", {"latent_dim": 16}),
                                         ("MTNet", {"time_step": 3, "long_num": 2, "ar_size": 2, "cnn_hid_size": 8, "rnn_hid_sizes": [8]}),
                                         ("TCN", {"hidden_units": 16, "levels": 2})])
def test_automl_models_train_on_gpu(model, extra):
    from zoo.automl.model.time_sequence import TimeSequenceModel

    torch.manual_seed(0)
    np.random.seed(0)
    rng = np.random.RandomState(0)
    x = rng.randn(512, 9, 3).astype(np.float32)
    fsl = 2 if model == "Seq2seq" else 1
    y = np.stack([x[:, -1, 0] * 0.5 + x[:, -2, 0] * 0.3] * fsl, 1).astype(np.float32)
    m = TimeSequenceModel(future_seq_len=fsl)
    cfg = dict(model=model, lr=0.01, batch_size=128, epochs=3, dropout=0.0, dropout_1=0.0, dropout_2=0.0, cnn_dropout=0.0, rnn_dropout=0.0, **extra)
    first = m.fit_eval(x, y if fsl > 1 else y[:, 0], **cfg)
    for _ in range(5):
        last = m.fit_eval(x, y if fsl > 1 else y[:, 0], **cfg)
    assert next(m.model.model.parameters()).is_cuda and last < first
    mean, std = m.predict_with_uncertainty(x[:8], n_iter=3)
    assert mean.shape == (8, fsl)


def test_nnframes_on_gpu():
    from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
    from zoo.pipeline.api.keras.optimizers import Adam
    from zoo.pipeline.nnframes import NNClassifier

    torch.manual_seed(0)
    rng = np.random.RandomState(0)
    x = rng.randn(800, 8).astype(np.float32)
    df = pd.DataFrame({"features": list(x), "label": (x.sum(1) > 0).astype(np.float32) + 1.0})
    from zoo.ops.dense import FusedDense

    model = nn.Sequential(FusedDense(8, 32, activation="relu"), FusedDense(32, 2), nn.Softmax(-1))
    m = NNClassifier(model, SparseCategoricalCrossEntropy(zero_based_label=False), [8]).setBatchSize(100).setMaxEpoch(15).setOptimMethod(Adam(lr=1e-2)).fit(df)
    assert (m.transform(df)["prediction"] == df["label"]).mean() > 0.9
