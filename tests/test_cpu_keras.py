"""Keras-style API on CPU: layer shapes and numerics against torch, Sequential / functional Model, autograd ops,
CustomLoss, Lambda, keras2 names (reference strategy: per-layer parity specs)."""
import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F


@pytest.fixture(autouse=True)
def _cpu_ctx(monkeypatch):
    import zoo.common.nncontext as nc

    monkeypatch.setenv("ZOO_FORCE_CPU", "1")
    nc._CONTEXT = None
    yield
    nc._CONTEXT = None


import zoo.pipeline.api.autograd as A
from zoo.pipeline.api.keras import layers as L
from zoo.pipeline.api.keras.engine.topology import Model, Sequential

CASES = [
    (lambda: L.Dense(5, activation="relu", input_shape=(7,)), (7,), (5,)),
    (lambda: L.Dense(5, input_shape=(3, 7)), (3, 7), (3, 5)),
    (lambda: L.Activation("tanh", input_shape=(4,)), (4,), (4,)),
    (lambda: L.Dropout(0.3, input_shape=(4,)), (4,), (4,)),
    (lambda: L.SpatialDropout1D(0.3, input_shape=(4, 5)), (4, 5), (4, 5)),
    (lambda: L.SpatialDropout2D(0.3, input_shape=(3, 4, 5)), (3, 4, 5), (3, 4, 5)),
    (lambda: L.Reshape((2, 6), input_shape=(3, 4)), (3, 4), (2, 6)),
    (lambda: L.Reshape((-1, 2), input_shape=(3, 4)), (3, 4), (6, 2)),
    (lambda: L.Permute((2, 1), input_shape=(3, 4)), (3, 4), (4, 3)),
    (lambda: L.Flatten(input_shape=(3, 4, 2)), (3, 4, 2), (24,)),
    (lambda: L.RepeatVector(3, input_shape=(5,)), (5,), (3, 5)),
    (lambda: L.Masking(0.0, input_shape=(3, 4)), (3, 4), (3, 4)),
    (lambda: L.MaxoutDense(4, nb_feature=3, input_shape=(6,)), (6,), (4,)),
    (lambda: L.Highway(activation="relu", input_shape=(6,)), (6,), (6,)),
    (lambda: L.Convolution1D(8, 3, input_shape=(10, 4)), (10, 4), (8, 8)),
    (lambda: L.Convolution1D(8, 3, border_mode="same", input_shape=(10, 4)), (10, 4), (10, 8)),
    (lambda: L.AtrousConvolution1D(8, 3, atrous_rate=2, input_shape=(10, 4)), (10, 4), (6, 8)),
    (lambda: L.Convolution2D(6, 3, 3, input_shape=(3, 12, 12)), (3, 12, 12), (6, 10, 10)),
    (lambda: L.Convolution2D(6, 3, 3, border_mode="same", subsample=(2, 2), input_shape=(3, 12, 12)), (3, 12, 12), (6, 6, 6)),
    (lambda: L.Convolution2D(6, 3, 3, dim_ordering="tf", input_shape=(12, 12, 3)), (12, 12, 3), (10, 10, 6)),
    (lambda: L.AtrousConvolution2D(4, 3, 3, atrous_rate=(2, 2), input_shape=(3, 12, 12)), (3, 12, 12), (4, 8, 8)),
    (lambda: L.Convolution3D(4, 3, 3, 3, input_shape=(2, 8, 8, 8)), (2, 8, 8, 8), (4, 6, 6, 6)),
    (lambda: L.Deconvolution2D(3, 3, 3, input_shape=(4, 6, 6)), (4, 6, 6), (3, 8, 8)),
    (lambda: L.SeparableConvolution2D(8, 3, 3, input_shape=(4, 9, 9)), (4, 9, 9), (8, 7, 7)),
    (lambda: L.ShareConvolution2D(5, 3, 3, input_shape=(3, 9, 9)), (3, 9, 9), (5, 7, 7)),
    (lambda: L.UpSampling1D(2, input_shape=(3, 4)), (3, 4), (6, 4)),
    (lambda: L.UpSampling2D((2, 3), input_shape=(2, 3, 4)), (2, 3, 4), (2, 6, 12)),
    (lambda: L.UpSampling3D((2, 2, 2), input_shape=(2, 3, 4, 5)), (2, 3, 4, 5), (2, 6, 8, 10)),
    (lambda: L.ZeroPadding1D(2, input_shape=(3, 4)), (3, 4), (7, 4)),
    (lambda: L.ZeroPadding2D((1, 2), input_shape=(2, 3, 4)), (2, 3, 4), (2, 5, 8)),
    (lambda: L.ZeroPadding3D((1, 1, 1), input_shape=(2, 3, 4, 5)), (2, 3, 4, 5), (2, 5, 6, 7)),
    (lambda: L.Cropping1D((1, 1), input_shape=(5, 4)), (5, 4), (3, 4)),
    (lambda: L.Cropping2D(((1, 1), (1, 0)), input_shape=(2, 5, 6)), (2, 5, 6), (2, 3, 5)),
    (lambda: L.Cropping3D(((1, 1), (1, 1), (1, 1)), input_shape=(2, 5, 6, 7)), (2, 5, 6, 7), (2, 3, 4, 5)),
    (lambda: L.MaxPooling1D(2, input_shape=(8, 3)), (8, 3), (4, 3)),
    (lambda: L.AveragePooling1D(3, stride=2, input_shape=(9, 3)), (9, 3), (4, 3)),
    (lambda: L.MaxPooling2D((2, 2), input_shape=(3, 8, 8)), (3, 8, 8), (3, 4, 4)),
    (lambda: L.MaxPooling2D((3, 3), strides=(2, 2), border_mode="same", input_shape=(3, 9, 9)), (3, 9, 9), (3, 5, 5)),
    (lambda: L.AveragePooling2D((2, 2), input_shape=(3, 8, 8)), (3, 8, 8), (3, 4, 4)),
    (lambda: L.MaxPooling3D(input_shape=(2, 4, 4, 4)), (2, 4, 4, 4), (2, 2, 2, 2)),
    (lambda: L.AveragePooling3D(input_shape=(2, 4, 4, 4)), (2, 4, 4, 4), (2, 2, 2, 2)),
    (lambda: L.GlobalMaxPooling1D(input_shape=(5, 3)), (5, 3), (3,)),
    (lambda: L.GlobalAveragePooling1D(input_shape=(5, 3)), (5, 3), (3,)),
    (lambda: L.GlobalMaxPooling2D(input_shape=(3, 5, 5)), (3, 5, 5), (3,)),
    (lambda: L.GlobalAveragePooling2D(input_shape=(3, 5, 5)), (3, 5, 5), (3,)),
    (lambda: L.GlobalMaxPooling3D(input_shape=(3, 4, 4, 4)), (3, 4, 4, 4), (3,)),
    (lambda: L.GlobalAveragePooling3D(input_shape=(3, 4, 4, 4)), (3, 4, 4, 4), (3,)),
    (lambda: L.LSTM(6, input_shape=(5, 3)), (5, 3), (6,)),
    (lambda: L.LSTM(6, return_sequences=True, input_shape=(5, 3)), (5, 3), (5, 6)),
    (lambda: L.GRU(6, return_sequences=True, go_backwards=True, input_shape=(5, 3)), (5, 3), (5, 6)),
    (lambda: L.SimpleRNN(6, input_shape=(5, 3)), (5, 3), (6,)),
    (lambda: L.Bidirectional(L.LSTM(4, return_sequences=True), input_shape=(5, 3)), (5, 3), (5, 8)),
    (lambda: L.Bidirectional(L.GRU(4), merge_mode="sum", input_shape=(5, 3)), (5, 3), (4,)),
    (lambda: L.TimeDistributed(L.Dense(7), input_shape=(5, 3)), (5, 3), (5, 7)),
    (lambda: L.ConvLSTM2D(4, 3, input_shape=(3, 2, 6, 6)), (3, 2, 6, 6), (4, 6, 6)),
    (lambda: L.ConvLSTM2D(4, 3, return_sequences=True, input_shape=(3, 2, 6, 6)), (3, 2, 6, 6), (3, 4, 6, 6)),
    (lambda: L.ConvLSTM3D(2, 3, input_shape=(2, 1, 4, 4, 4)), (2, 1, 4, 4, 4), (2, 4, 4, 4)),
    (lambda: L.BatchNormalization(input_shape=(3, 4, 4)), (3, 4, 4), (3, 4, 4)),
    (lambda: L.BatchNormalization(input_shape=(6,)), (6,), (6,)),
    (lambda: L.LayerNorm(input_shape=(4, 6)), (4, 6), (4, 6)),
    (lambda: L.GaussianNoise(0.1, input_shape=(4,)), (4,), (4,)),
    (lambda: L.GaussianDropout(0.2, input_shape=(4,)), (4,), (4,)),
    (lambda: L.LeakyReLU(0.1, input_shape=(4,)), (4,), (4,)),
    (lambda: L.ELU(1.0, input_shape=(4,)), (4,), (4,)),
    (lambda: L.ThresholdedReLU(0.5, input_shape=(4,)), (4,), (4,)),
    (lambda: L.SReLU(input_shape=(4, 5)), (4, 5), (4, 5)),
    (lambda: L.LocallyConnected1D(6, 3, input_shape=(8, 4)), (8, 4), (6, 6)),
    (lambda: L.LocallyConnected2D(5, 3, 3, input_shape=(2, 7, 7)), (2, 7, 7), (5, 5, 5)),
    (lambda: L.Select(1, 2, input_shape=(4, 5)), (4, 5), (5,)),
    (lambda: L.Narrow(1, 1, 2, input_shape=(4, 5)), (4, 5), (2, 5)),
    (lambda: L.Squeeze(1, input_shape=(1, 5)), (1, 5), (5,)),
    (lambda: L.ExpandDim(1, input_shape=(5,)), (5,), (1, 5)),
    (lambda: L.AddConstant(1.5, input_shape=(5,)), (5,), (5,)),
    (lambda: L.MulConstant(2.0, input_shape=(5,)), (5,), (5,)),
    (lambda: L.LRN2D(input_shape=(6, 5, 5)), (6, 5, 5), (6, 5, 5)),
    (lambda: L.WithinChannelLRN2D(input_shape=(3, 5, 5)), (3, 5, 5), (3, 5, 5)),
    (lambda: L.CAdd((1, 5), input_shape=(5,)), (5,), (5,)),
    (lambda: L.CMul((1, 5), input_shape=(5,)), (5,), (5,)),
    (lambda: L.Mul(input_shape=(5,)), (5,), (5,)),
    (lambda: L.Scale((1, 5), input_shape=(5,)), (5,), (5,)),
    (lambda: L.Power(2.0, 2.0, 1.0, input_shape=(5,)), (5,), (5,)),
    (lambda: L.HardShrink(0.5, input_shape=(5,)), (5,), (5,)),
    (lambda: L.SoftShrink(0.5, input_shape=(5,)), (5,), (5,)),
    (lambda: L.HardTanh(-2, 2, input_shape=(5,)), (5,), (5,)),
    (lambda: L.PReLU(input_shape=(5,)), (5,), (5,)),
    (lambda: L.RReLU(input_shape=(5,)), (5,), (5,)),
    (lambda: L.BinaryThreshold(0.1, input_shape=(5,)), (5,), (5,)),
    (lambda: L.Threshold(0.1, 9.0, input_shape=(5,)), (5,), (5,)),
    (lambda: L.ResizeBilinear(6, 8, input_shape=(3, 4, 4)), (3, 4, 4), (3, 6, 8)),
    (lambda: L.Max(1, input_shape=(4, 5)), (4, 5), (4,)),
    (lambda: L.Exp(input_shape=(3,)), (3,), (3,)),
    (lambda: L.Square(input_shape=(3,)), (3,), (3,)),
    (lambda: L.Negative(input_shape=(3,)), (3,), (3,)),
    (lambda: L.Identity(input_shape=(3,)), (3,), (3,)),
    (lambda: L.Softmax(input_shape=(3,)), (3,), (3,)),
]


@pytest.mark.parametrize("idx", range(len(CASES)))
def test_layer_shapes_and_gradients(idx):
    make, in_shape, out_shape = CASES[idx]
    torch.manual_seed(idx)
    layer = make()
    m = Sequential().add(layer)
    assert m.get_output_shape() == (None,) + out_shape, type(layer).__name__
    x = torch.rand(4, *in_shape) + 0.1
    x.requires_grad_(True)
    y = m(x)
    assert tuple(y.shape) == (4,) + out_shape
    assert torch.isfinite(y).all()
    if y.requires_grad:
        y.float().sum().backward()
    assert layer.get_input_shape() == (None,) + in_shape


def test_selected_numerics():
    torch.manual_seed(0)
    d = L.Dense(4, activation="sigmoid", input_shape=(6,))
    m = Sequential([d])
    x = torch.randn(5, 6)
    torch.testing.assert_close(m(x), torch.sigmoid(F.linear(x, d.dense.weight, d.dense.bias)))
    assert m.logits_kind() == "sigmoid"
    torch.testing.assert_close(m.forward_logits(x), F.linear(x, d.dense.weight, d.dense.bias))
    c = L.Convolution1D(3, 2, input_shape=(5, 4))
    Sequential([c])
    x = torch.randn(2, 5, 4)
    torch.testing.assert_close(c(x), c.conv(x.transpose(1, 2)).transpose(1, 2))
    mp = L.MaxPooling2D((3, 3), strides=(2, 2), border_mode="same", input_shape=(1, 5, 5))
    x = torch.arange(25.0).reshape(1, 1, 5, 5)
    assert Sequential([mp])(x)[0, 0].tolist() == [[6, 8, 9], [16, 18, 19], [21, 23, 24]]
    e = L.Embedding(10, 3, input_shape=(4,), zero_based_id=False)
    Sequential([e])
    ids = torch.tensor([[1, 2, 3, 10]])
    torch.testing.assert_close(e(ids), e.table.weight[ids - 1])
    se = L.SparseEmbedding(10, 3, combiner="mean", input_shape=(4,))
    Sequential([se])
    ids = torch.tensor([[0, 2, -1, -1], [5, -1, -1, -1]])
    w = se.table.weight
    torch.testing.assert_close(se(ids), torch.stack([(w[0] + w[2]) / 2, w[5]]))
    mk = L.Masking(0.0, input_shape=(3, 2))
    x = torch.tensor([[[1.0, 0.0], [0.0, 0.0], [0.0, 2.0]]])
    assert mk(x).tolist() == x.tolist()
    lstm = L.LSTM(5, inner_activation="sigmoid", return_sequences=True, input_shape=(4, 3))
    Sequential([lstm])
    x = torch.randn(2, 4, 3)
    torch.testing.assert_close(lstm(x), lstm.cell.rnn(x)[0])


def test_merge_modes_and_functional_model():
    a, b = torch.randn(3, 4), torch.randn(3, 4)
    for mode, ref in [("sum", a + b), ("mul", a * b), ("ave", (a + b) / 2), ("max", torch.maximum(a, b)), ("min", torch.minimum(a, b)),
                      ("sub", a - b), ("div", a / b), ("concat", torch.cat([a, b], -1)), ("dot", (a * b).sum(1, keepdim=True))]:
        torch.testing.assert_close(L.Merge(mode=mode)(a, b), ref)
    assert L.Merge(mode="cos")(a, b).shape == (3, 1, 1)
    with pytest.raises(ValueError):
        L.Merge(mode="bogus")
    # functional: two inputs, shared layer, merge, autograd ops
    x1, x2 = L.Input(shape=(8,)), L.Input(shape=(8,))
    shared = L.Dense(6, activation="relu")
    h = L.merge([shared(x1), shared(x2)], mode="concat")
    out = L.Dense(2, activation="softmax")(h)
    model = Model([x1, x2], out)
    assert model.get_output_shape() == (None, 2) and len(list(shared.parameters())) == 2
    n_params = sum(p.numel() for p in model.parameters())
    assert n_params == (8 * 6 + 6) + (12 * 2 + 2)                # the shared Dense is counted once
    y = model(torch.randn(5, 8), torch.randn(5, 8))
    torch.testing.assert_close(y.sum(1), torch.ones(5))
    sub = model.new_graph([shared.zoo_name])
    assert sub(torch.randn(5, 8), torch.randn(5, 8)).shape == (5, 6)
    model.freeze_up_to([shared.zoo_name])
    assert not any(p.requires_grad for p in shared.parameters()) and any(p.requires_grad for p in model.parameters())


def test_autograd_ops_lambda_customloss_parameter():
    x = A.Variable(input_shape=(3, 4))
    y = A.Variable(input_shape=(3, 4))
    xt, yt = torch.rand(2, 3, 4) + 0.5, torch.rand(2, 3, 4) + 0.5
    run = lambda v, *ts: A.GraphModule([x, y][:len(ts)], v)(*ts)
    torch.testing.assert_close(run(A.mean(x, axis=1), xt), xt.mean(1))
    torch.testing.assert_close(run(A.sum(x, axis=2, keepDims=True), xt), xt.sum(2, keepdim=True))
    torch.testing.assert_close(run(A.clip(x, 0.6, 0.9), xt), xt.clamp(0.6, 0.9))
    torch.testing.assert_close(run(A.sqrt(A.square(x)) + A.exp(A.log(x)), xt), 2 * xt)
    torch.testing.assert_close(run(A.maximum(x, y), xt, yt), torch.maximum(xt, yt))
    torch.testing.assert_close(run(x * y - x / y + 2.0 - (-x), xt, yt), xt * yt - xt / yt + 2.0 + xt)
    torch.testing.assert_close(run(1.0 - x, xt), 1.0 - xt)
    torch.testing.assert_close(run(A.mm(x, y, axes=[2, 2]), xt, yt), xt @ yt.transpose(1, 2))
    torch.testing.assert_close(run(A.batch_dot(x, y, axes=[2, 2]), xt, yt), torch.bmm(xt, yt.transpose(1, 2)))
    torch.testing.assert_close(run(A.l2_normalize(x, 2), xt), F.normalize(xt, dim=2))
    torch.testing.assert_close(run(A.stack([x, y], axis=1), xt, yt), torch.stack([xt, yt], 1))
    torch.testing.assert_close(run(A.expand_dims(x, 1), xt), xt.unsqueeze(1))
    torch.testing.assert_close(run(x.slice(1, 1, 2), xt), xt[:, 1:3])
    torch.testing.assert_close(run(x.index_select(1, 0), xt), xt[:, 0])
    torch.testing.assert_close(run(A.erf(x), xt), torch.erf(xt))
    assert A.mean(x, axis=1).get_output_shape() == (None, 4)
    lam = Sequential([A.Lambda(lambda v: A.softsign(v) * 2.0, input_shape=(4,))])
    torch.testing.assert_close(lam(xt[:, 0]), F.softsign(xt[:, 0]) * 2)
    two = A.Lambda(lambda a, b: a + b)
    assert two(torch.ones(2, 3), torch.ones(2, 3)).sum() == 12
    loss = A.CustomLoss(lambda y_true, y_pred: A.mean(A.abs(y_true - y_pred), axis=1), y_pred_shape=(3,))
    yp, ytrue = torch.tensor([[1.0, 2.0, 3.0]]), torch.tensor([[2.0, 2.0, 5.0]])
    assert abs(float(loss(yp, ytrue)) - 1.0) < 1e-6
    assert loss.backward(ytrue.numpy(), yp.numpy()).shape == (1, 3)
    w = A.Parameter((1, 4), init_weight=np.ones((1, 4)))
    out = A.Variable(input_shape=(4,)) if False else None
    xin = A.Variable(input_shape=(4,))
    g = A.GraphModule([xin], xin * w + A.Constant(np.full((1, 4), 0.5)))
    torch.testing.assert_close(g(torch.ones(3, 4)), torch.full((3, 4), 1.5))
    assert len(list(g.parameters())) == 1 and w.get_weight().shape == (1, 4)


def test_transformer_and_bert():
    torch.manual_seed(0)
    t = L.TransformerLayer.init(vocab=50, seq_len=6, n_block=2, n_head=2, hidden_size=8, hidden_drop=0.0, attn_drop=0.0, embedding_drop=0.0)
    ids = torch.randint(0, 44, (3, 6))
    pos = torch.arange(44, 50).repeat(3, 1)
    h, pooled = t(ids, pos)
    assert h.shape == (3, 6, 8) and pooled.shape == (3, 8)
    ids2 = ids.clone()
    ids2[:, -1] = (ids2[:, -1] + 1) % 44                     # causal: earlier positions must not change
    h2, _ = t.eval()(ids2, pos)
    h1, _ = t(ids, pos)
    torch.testing.assert_close(h1[:, :-1], h2[:, :-1])
    assert not torch.allclose(h1[:, -1], h2[:, -1])
    b = L.BERT.init(vocab=30, hidden_size=8, n_block=2, n_head=2, seq_len=5, intermediate_size=16, hidden_drop=0.0, attn_drop=0.0)
    ids = torch.randint(0, 30, (2, 5))
    outs = b.eval()(ids, torch.zeros(2, 5, dtype=torch.long), torch.arange(5).repeat(2, 1), torch.ones(2, 1, 1, 5))
    assert len(outs) == 3 and outs[0].shape == (2, 5, 8) and outs[-1].shape == (2, 8)
    mask = torch.ones(2, 1, 1, 5)
    mask[:, :, :, -1] = 0                                      # masking the last token: changing it leaves the others unchanged
    ids2 = ids.clone()
    ids2[:, -1] = (ids2[:, -1] + 1) % 30
    o1 = b(ids, torch.zeros(2, 5, dtype=torch.long), torch.arange(5).repeat(2, 1), mask)
    o2 = b(ids2, torch.zeros(2, 5, dtype=torch.long), torch.arange(5).repeat(2, 1), mask)
    torch.testing.assert_close(o1[-2][:, :-1], o2[-2][:, :-1], atol=1e-5, rtol=1e-4)
    with pytest.raises(TypeError):
        L.BERT(1, 1, 4, 0, 0, 0.02, True, None, input_shape=((5,), (5,)))


def test_sequential_fit_and_keras2(tmp_path):
    from zoo.pipeline.api.keras2 import layers as K2

    torch.manual_seed(0)
    rng = np.random.RandomState(0)
    x = rng.randn(600, 10).astype(np.float32)
    y = (x[:, :5].sum(1) > x[:, 5:].sum(1)).astype(np.int64)
    m = Sequential()
    m.add(K2.Dense(16, activation="relu", input_shape=(10,))).add(K2.Dropout(0.1)).add(K2.Dense(2, activation="softmax"))
    m.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    m.fit(x, y, batch_size=60, nb_epoch=30)
    loss, acc = m.evaluate(x, y, batch_size=200)
    assert acc > 0.9
    assert m.predict_classes(x[:5]).shape == (5,)
    m.summary()
    with pytest.raises(ValueError):
        Sequential().add(L.Dense(3))
    c = Sequential([K2.Conv2D(4, 3, input_shape=(1, 8, 8)), K2.MaxPooling2D(), K2.Flatten(), K2.Dense(3)])
    assert c.get_output_shape() == (None, 3)
    a, b = L.Input(shape=(4,)), L.Input(shape=(4,))
    mm = Model([a, b], K2.maximum([a, b]))
    torch.testing.assert_close(mm(torch.zeros(2, 4), torch.ones(2, 4)), torch.ones(2, 4))


@pytest.mark.parametrize("idx", range(len(CASES)))
def test_layer_serializer_roundtrip(idx, tmp_path):
    """Every layer survives saveModel -> Net.load with identical outputs (reference: the serializer specs, one per layer)."""
    from zoo.pipeline.api.net import Net

    make, in_shape, _ = CASES[idx]
    torch.manual_seed(100 + idx)
    m = Sequential().add(make())
    m.eval()
    x = torch.rand(3, *in_shape) + 0.1
    with torch.no_grad():
        want = m(x)
    path = str(tmp_path / "layer.model")
    m.saveModel(path, over_write=True)
    m2 = Net.load(path)
    m2.eval()
    with torch.no_grad():
        got = m2(x)
    torch.testing.assert_close(got, want)
    assert [tuple(p.shape) for p in m2.parameters()] == [tuple(p.shape) for p in m.parameters()]


@pytest.mark.parametrize("L,causal,masked", [(77, False, False), (300, True, False), (300, False, True), (130, True, True)])
def test_blockwise_attention_restatement_matches_dense_attention(L, causal, masked):
    """The algorithm of the K8 kernel (online softmax over key blocks, bf16 probabilities) against the reference's dense attention."""
    from zoo.ops.attention import attention_blockwise_reference, attention_reference

    torch.manual_seed(L)
    q, k, v = (torch.randn(2, 2, L, 64).to(torch.bfloat16) for _ in range(3))
    km = None
    if masked:
        keep = torch.rand(2, L) > 0.2
        keep[:, 0] = True
        km = (1.0 - keep.float()) * -10000.0
    got = attention_blockwise_reference(q, k, v, km, causal)
    want = attention_reference(q.float(), k.float(), v.float(), None if km is None else km[:, None, None, :], causal)
    torch.testing.assert_close(got, want, atol=1e-2, rtol=1e-2)


def test_save_to_tf_and_save_to_keras2(tmp_path):
    """KerasNet.saveToTf / saveToKeras2 (Topology.scala:559-575; NetSpec.scala:147-192): the TF export is read back by TFNet and gives
    the model's output; the Keras-2 export is a weights file in Keras layout plus the script that rebuilds the model."""
    from zoo.pipeline.api.keras.layers import LSTM, Dense, Dropout, Permute, Reshape
    from zoo.pipeline.api.keras.models import Sequential
    from zoo.pipeline.api.net import TFNet
    from zoo.util.hdf5 import File

    torch.manual_seed(0)
    m = Sequential()
    m.add(Dense(8, activation="tanh", input_shape=(6,)))
    m.add(Dropout(0.5))
    m.add(Dense(3, activation="softmax"))
    m.eval()
    x = torch.randn(4, 6)
    ins, outs = m.saveToTf(str(tmp_path / "tf"))
    back = TFNet.from_export_folder(str(tmp_path / "tf"))
    assert torch.allclose(back(x), m(x), atol=1e-5)

    g = Sequential()
    g.add(Reshape((13, 4), input_shape=(52,)))
    g.add(Permute((2, 1)))
    g.add(LSTM(10, return_sequences=True))
    g.add(LSTM(10))
    g.add(Dropout(0.5))
    g.add(Dense(2, activation="softmax"))
    weights, script = g.saveToKeras2(str(tmp_path / "model.h5"))
    code = open(script).read()
    compile(code, script, "exec")                                   # valid Python
    assert "Reshape((13, 4)" in code and "Permute((2, 1)" in code and code.count("LSTM(10") == 2 and "recurrent_activation='hard_sigmoid'" in code
    assert "Dense(2, activation='softmax'" in code and "input_shape=(52,)" in code and "model.load_weights" in code
    with File(weights) as f:
        layers = [n.decode() for n in f.attrs["layer_names"]]
        assert len(layers) == 3                                     # two LSTMs and the Dense carry weights
        dense = layers[-1]
        kernel = next(np.asarray(f[dense][w.decode()]) for w in f[dense].attrs["weight_names"] if np.asarray(f[dense][w.decode()]).ndim == 2)
        assert kernel.shape == (10, 2)                              # [in, out], Keras' layout
    with pytest.raises(NotImplementedError):
        from zoo.pipeline.api.keras.layers import Convolution2D

        c = Sequential()
        c.add(Convolution2D(2, 3, 3, input_shape=(1, 8, 8)))
        c.saveToTf(str(tmp_path / "c"))
