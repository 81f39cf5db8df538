"""Other built-in models and the image / image3d / text feature pipelines on CPU
(reference tests: pyzoo/test/zoo/models/**, pyzoo/test/zoo/feature/**)."""
import os

import numpy as np
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


def test_anomaly_detector(tmp_path):
    from zoo.models.anomalydetection import AnomalyDetector

    torch.manual_seed(0)
    t = np.arange(400, dtype=np.float32)
    series = np.sin(t / 8.0)
    unrolled = AnomalyDetector.unroll(series[:, None], unroll_length=10, predict_step=1)
    assert len(unrolled) == 390 and unrolled[0].feature.shape == (10, 1) and unrolled[0].label == pytest.approx(series[10]) and unrolled[5].index == 5
    train, test = AnomalyDetector.train_test_split(unrolled, 50)
    x, y = AnomalyDetector.to_arrays(train)
    m = AnomalyDetector(feature_shape=(10, 1), hidden_layers=[8, 16, 8], dropouts=[0.0, 0.0, 0.0])
    m.compile(optimizer="adam", loss="mse")
    before = m.evaluate(x, y, batch_size=128)[0]
    m.fit(x, y, batch_size=64, nb_epoch=8)
    assert m.evaluate(x, y, batch_size=128)[0] < before
    xt, yt = AnomalyDetector.to_arrays(test)
    yp = m.predict(xt)
    yt2 = yt.copy()
    yt2[7] += 5.0
    an = AnomalyDetector.detect_anomalies(yt2, yp, anomaly_size=1)
    assert [i for i, a in enumerate(an) if a[2] is not None] == [7]
    m.save_model(str(tmp_path / "ad.model"), over_write=True)
    m2 = AnomalyDetector.load_model(str(tmp_path / "ad.model"))
    np.testing.assert_allclose(m2.predict(xt), yp, atol=1e-6)
    with pytest.raises(TypeError):
        AnomalyDetector((10, 1), bogus=1)
    with pytest.raises(AssertionError):
        AnomalyDetector((10, 1), hidden_layers=[8, 8], dropouts=[0.1])


def _glove(tmp_path, words, dim=8):
    rng = np.random.RandomState(0)
    p = tmp_path / "glove.txt"
    with open(p, "w") as f:
        for w in words:
            f.write(w + " " + " ".join(f"{v:.4f}" for v in rng.randn(dim)) + "\n")
    return str(p)


@pytest.mark.parametrize("encoder", ["cnn", "lstm", "gru"])
def test_text_classifier(tmp_path, encoder):
    from zoo.models.textclassification import TextClassifier
    from zoo.pipeline.api.keras.layers import WordEmbedding

    words = [f"w{i}" for i in range(30)]
    path = _glove(tmp_path, words)
    wi = WordEmbedding.get_word_index(path)
    assert wi["w0"] == 1 and len(wi) == 30
    m = TextClassifier(3, path, wi, sequence_length=12, encoder=encoder, encoder_output_dim=16)
    x = np.random.randint(1, 31, (40, 12))
    y = (x[:, 0] % 3).astype(np.int64)
    out = m(torch.from_numpy(x))
    assert out.shape == (40, 3) and torch.allclose(out.sum(1), torch.ones(40), atol=1e-5)
    m.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    m.fit(x, y, batch_size=20, nb_epoch=2)
    assert m.predict(x).shape == (40, 3)
    assert not m.embedding.weight.requires_grad                       # GloVe vectors stay frozen
    with pytest.raises(ValueError):
        TextClassifier(3, path, wi, encoder="bogus")
    t = TextClassifier(2, vocab_size=50, embed_dim=8, sequence_length=6)
    assert t(torch.randint(0, 50, (4, 6))).shape == (4, 2)


def test_text_set_pipeline_and_knrm(tmp_path):
    from zoo.feature.common import Relation
    from zoo.feature.text import LocalTextSet, TextSet
    from zoo.models.textmatching import KNRM

    ts = LocalTextSet(["Hello, World! hello zoo", "the quick brown fox the end", "zoo zoo zoo"], [0, 1, 1])
    tok = ts.tokenize().normalize()
    assert tok.features[0]["tokens"] == ["hello", "world", "hello", "zoo"]
    idx = tok.word2idx()
    wi = idx.get_word_index()
    assert wi["zoo"] == 1 and wi["hello"] == 2 and wi["the"] == 3          # by descending frequency, then alphabetical
    shaped = idx.shape_sequence(5)
    assert shaped.features[0]["indexedTokens"] == [2.0, float(wi["world"]), 2.0, 1.0, 0.0]
    assert shaped.features[1]["indexedTokens"][0] == float(wi["quick"])     # 'pre' truncation keeps the LAST 5 of 6
    post = idx.shape_sequence(5, trunc_mode="post")
    assert post.features[1]["indexedTokens"][0] == float(wi["the"])
    samples = shaped.generate_sample().get_samples()
    assert samples[0].feature.shape == (5,) and float(samples[2].label[0]) == 1.0
    idx.save_word_index(str(tmp_path / "wi.txt"))
    again = TextSet().load_word_index(str(tmp_path / "wi.txt")).get_word_index()
    assert again == wi
    limited = ts.tokenize().normalize().word2idx(remove_topN=1, max_words_num=2).get_word_index()
    assert "zoo" not in limited and len(limited) == 2
    x, y = shaped.to_arrays()
    assert x.shape == (3, 5) and y.tolist() == [0, 1, 1]
    # directory / csv readers
    os.makedirs(tmp_path / "news" / "a"); os.makedirs(tmp_path / "news" / "b")
    (tmp_path / "news" / "a" / "1.txt").write_text("alpha text"); (tmp_path / "news" / "b" / "2.txt").write_text("beta text")
    rd = TextSet.read(str(tmp_path / "news"))
    assert rd.get_labels() == [0, 1] and rd.get_texts() == ["alpha text", "beta text"]
    (tmp_path / "q.csv").write_text("q1,what is zoo\nq2,quick fox\n")
    (tmp_path / "d.csv").write_text("d1,zoo is a framework\nd2,the fox is quick\nd3,nothing here\n")
    q = TextSet.read_csv(str(tmp_path / "q.csv")).tokenize().normalize().word2idx().shape_sequence(4)
    d = TextSet.read_csv(str(tmp_path / "d.csv")).tokenize().normalize().word2idx(existing_map=q.get_word_index()).shape_sequence(6)
    rels = [Relation("q1", "d1", 1), Relation("q1", "d3", 0), Relation("q2", "d2", 1), Relation("q2", "d1", 0), Relation("q2", "d3", 0)]
    pairs = TextSet.from_relation_pairs(rels, q, d)
    assert len(pairs) == 3 and pairs.get_samples()[0].feature.shape == (2, 10) and pairs.get_samples()[0].label.tolist() == [1.0, 0.0]
    lists = TextSet.from_relation_lists(rels, q, d)
    assert len(lists) == 2 and lists.get_samples()[1].feature.shape == (3, 10) and lists.get_samples()[1].label.shape == (3, 1)
    vocab = max(d.get_word_index().values()) + 1
    knrm = KNRM(4, 6, vocab_size=vocab, embed_size=8, kernel_num=5)
    out = knrm(torch.from_numpy(lists.get_samples()[1].feature).long())
    assert out.shape == (3, 1)
    assert 0.0 <= knrm.evaluate_ndcg(lists, 3) <= 1.0 and 0.0 <= knrm.evaluate_map(lists) <= 1.0
    assert torch.allclose(knrm.mu.reshape(-1), torch.tensor([-0.75, -0.25, 0.25, 0.75, 1.0])) and float(knrm.sig.reshape(-1)[-1]) == pytest.approx(0.001)
    cls_ = KNRM(4, 6, vocab_size=vocab, embed_size=8, target_mode="classification")
    o = cls_(torch.randint(0, vocab, (5, 10)))
    assert ((o >= 0) & (o <= 1)).all()
    with pytest.raises(ValueError):
        KNRM(4, 6, vocab_size=vocab, embed_size=8, target_mode="bogus")


def test_seq2seq_fit_and_infer():
    from zoo.models.seq2seq import Bridge, RNNDecoder, RNNEncoder, Seq2seq

    torch.manual_seed(0)
    enc = RNNEncoder.initialize("lstm", 1, 16, input_shape=(5, 3))
    dec = RNNDecoder.initialize("lstm", 1, 16, input_shape=(4, 3))
    model = Seq2seq(enc, dec, [5, 3], [4, 3], bridge=Bridge.initialize("dense", 16), generator=nn.Linear(16, 3))
    rng = np.random.RandomState(0)
    a = rng.randn(128, 5, 3).astype(np.float32)
    b_in = np.concatenate([np.zeros((128, 1, 3), np.float32), np.repeat(a[:, -1:, :], 3, 1)], 1)
    tgt = np.repeat(a[:, -1:, :], 4, 1)                        # copy the last encoder step 4 times
    model.compile(optimizer="adam", loss="mse")
    l0 = model.evaluate([a, b_in], tgt, batch_size=64)[0]
    model.fit([a, b_in], tgt, batch_size=32, nb_epoch=30)
    assert model.evaluate([a, b_in], tgt, batch_size=64)[0] < l0 * 0.8
    out = model.infer(a[0], start_sign=np.zeros(3, np.float32), max_seq_len=4)
    assert out.shape == (1, 4, 3)
    g = Seq2seq(RNNEncoder.initialize("gru", 2, 8, input_shape=(5, 3)), RNNDecoder.initialize("gru", 2, 8, input_shape=(4, 3)), [5, 3], [4, 3])
    assert g(torch.randn(2, 5, 3), torch.randn(2, 4, 3)).shape == (2, 4, 8)
    with pytest.raises(ValueError):
        RNNEncoder.initialize("bogus", 1, 4)
    with pytest.raises(ValueError):
        Bridge.initialize("bogus", 4)


def test_image_set_and_preprocessing(tmp_path):
    import cv2

    from zoo.feature.image import (ChainedPreprocessing, ImageBrightness, ImageBytesToMat, ImageCenterCrop, ImageChannelNormalize, ImageColorJitter,
                                   ImageExpand, ImageFiller, ImageFixedCrop, ImageHFlip, ImageMatToTensor, ImagePixelNormalize, ImageRandomCrop,
                                   ImageRandomPreprocessing, ImageResize, ImageSet, ImageSetToSample, ImageAspectScale, PerImageNormalize, ImageHue,
                                   ImageSaturation, ImageChannelOrder)

    rng = np.random.RandomState(0)
    for c in ("cats", "dogs"):
        os.makedirs(tmp_path / c)
        for i in range(3):
            cv2.imwrite(str(tmp_path / c / f"{i}.png"), rng.randint(0, 256, (40 + i, 50, 3), dtype=np.uint8))
    s = ImageSet.read(str(tmp_path), with_label=True)
    assert len(s) == 6 and s.label_map == {"cats": 1, "dogs": 2} and [float(l[0]) for l in s.get_label()] == [1, 1, 1, 2, 2, 2]
    assert ImageSet.read(str(tmp_path), resize_height=20, resize_width=30).features[0]["mat"].shape == (20, 30, 3)
    chain = ChainedPreprocessing([ImageResize(32, 32), ImageCenterCrop(24, 24), ImageChannelNormalize(100.0, 110.0, 120.0, 2.0, 2.0, 2.0),
                                  ImageMatToTensor(), ImageSetToSample()])
    out = s.transform(chain)
    f0 = out.features[0]
    assert f0["imageTensor"].shape == (3, 24, 24) and f0["sample"].feature.shape == (3, 24, 24) and float(f0["sample"].label[0]) == 1.0
    ref = (cv2.resize(s.features[0]["mat"], (32, 32))[4:28, 4:28].astype(np.float32) - np.array([120.0, 110.0, 100.0])) / 2.0
    np.testing.assert_allclose(f0["imageTensor"], ref.transpose(2, 0, 1), atol=1e-5)
    assert s.features[0]["mat"].shape == (40, 50, 3)                   # the source set is untouched
    fs = out.to_feature_set()
    b = next(iter(fs.batches(4, train=False)))
    assert tuple(b.inputs.shape) == (4, 3, 24, 24) and b.targets.shape[0] == 4
    mat = s.features[0]["mat"]
    one = lambda op: op({"mat": mat.copy(), "originalSize": mat.shape, "uri": "x"})["mat"]
    assert np.array_equal(one(ImageHFlip()), mat[:, ::-1])
    assert one(ImageRandomCrop(20, 10)).shape == (10, 20, 3)
    assert one(ImageFixedCrop(0.1, 0.2, 0.5, 0.7)).shape == (20, 20, 3)
    assert (one(ImageFiller(0.0, 0.0, 0.5, 0.5, 7))[:20, :25] == 7).all()
    e = one(ImageExpand(min_expand_ratio=2.0, max_expand_ratio=2.0))
    assert e.shape == (80, 100, 3)
    assert one(ImageAspectScale(20, max_size=30)).shape[:2] == (20, 25)
    assert one(ImageAspectScale(20, max_size=20)).shape[:2] == (16, 20)        # the long side is capped
    pn = one(PerImageNormalize(0.0, 1.0))
    assert pn.min() == 0.0 and pn.max() == pytest.approx(1.0)
    assert one(ImagePixelNormalize(np.ones(mat.size))).dtype == np.float32
    for op in (ImageBrightness(-10, 10), ImageHue(-18, 18), ImageSaturation(0.5, 1.5), ImageChannelOrder(), ImageColorJitter(shuffle=True)):
        assert one(op).shape == mat.shape
    assert ImageRandomPreprocessing(ImageHFlip(), 0.0)({"mat": mat})["mat"] is mat
    raw = open(tmp_path / "cats" / "0.png", "rb").read()
    assert np.array_equal(ImageBytesToMat()({"bytes": raw})["mat"], mat)


def test_image3d_transforms():
    from zoo.feature.image3d import AffineTransform3D, CenterCrop3D, Crop3D, RandomCrop3D, Rotate3D

    v = np.arange(8 * 9 * 10, dtype=np.float32).reshape(8, 9, 10)
    assert np.array_equal(Crop3D([1, 2, 3], [4, 4, 4])(v), v[1:5, 2:6, 3:7])
    assert CenterCrop3D(4, 5, 6)(v).shape == (4, 5, 6) and RandomCrop3D(3, 3, 3)(v).shape == (3, 3, 3)
    with pytest.raises(ValueError):
        Crop3D([6, 0, 0], [4, 4, 4])(v)
    np.testing.assert_allclose(AffineTransform3D(np.eye(3))(v), v, atol=1e-4)
    shifted = AffineTransform3D(np.eye(3), translation=[0, 0, 1], clamp_mode="padding", pad_val=-1.0)(v)
    np.testing.assert_allclose(shifted[:, :, :-1], v[:, :, 1:], atol=1e-4)
    assert (shifted[:, :, -1] == -1.0).all()
    cube = np.random.RandomState(0).rand(6, 6, 6).astype(np.float32)
    r = Rotate3D([np.pi / 2, 0, 0])(cube)                              # 90 degrees about the depth axis... one of the planes
    assert r.shape == cube.shape and abs(float(r.sum()) - float(cube.sum())) < 1e-2
    np.testing.assert_allclose(Rotate3D([0, 0, 0])(cube), cube, atol=1e-5)
    f = {"imageTensor": v}
    assert CenterCrop3D(2, 2, 2)(f)["imageTensor"].shape == (2, 2, 2)


def test_image_models_and_detection_utils(tmp_path):
    import cv2

    from zoo.feature.image import ImageSet
    from zoo.models.image.imageclassification import ImageClassifier, LabelOutput
    from zoo.models.image.objectdetection import DecodeOutput, ObjectDetector, ScaleDetection, Visualizer, read_pascal_label_map
    from zoo.models.image.objectdetection.object_detector import nms

    rng = np.random.RandomState(0)
    os.makedirs(tmp_path / "imgs")
    for i in range(3):
        cv2.imwrite(str(tmp_path / "imgs" / f"{i}.jpg"), rng.randint(0, 256, (300, 280, 3), dtype=np.uint8))
    net = nn.Sequential(nn.Conv2d(3, 4, 3, stride=8), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(4, 10), nn.Softmax(-1))
    clf = ImageClassifier(net)
    out = clf.predict_image_set(ImageSet.read(str(tmp_path / "imgs")))
    f = out.features[0]
    assert f["predict"].shape == (10,) and len(f["clses"]) == 10 and f["probs"][0] >= f["probs"][1] and f["imageTensor"].shape == (3, 224, 224)
    torch.save(net, tmp_path / "net.pt")
    assert isinstance(ImageClassifier.load_model(str(tmp_path / "net.pt")), ImageClassifier)
    lo = LabelOutput({0: "a", 1: "b"})({"predict": np.array([0.2, 0.8])})
    assert lo["clses"] == ["b", "a"]
    assert nms(np.array([[0, 0, 10, 10], [1, 1, 10, 10], [20, 20, 30, 30]], np.float32), np.array([0.9, 0.8, 0.7]), 0.5) == [0, 2]
    feat = {"mat": np.zeros((100, 200, 3), np.uint8), "originalSize": (100, 200, 3),
            "predict": np.array([2, 15, 0.9, 0.1, 0.2, 0.5, 0.6, 7, 0.2, 0, 0, 1, 1], np.float32)}
    DecodeOutput()(feat)
    assert feat["predict"].shape == (2, 6)
    ScaleDetection()(feat)
    np.testing.assert_allclose(feat["predict"][0], [15, 0.9, 20, 20, 100, 60])
    Visualizer(read_pascal_label_map(), thresh=0.3)(feat)
    assert feat["visualized"][:4] == b"\x89PNG"
    assert isinstance(ObjectDetector(net).get_config().label_map(), dict)
