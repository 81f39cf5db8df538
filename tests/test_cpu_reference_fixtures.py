"""The data readers against the reference's own test resources: files written by Spark, Hadoop, TensorFlow and the reference
itself, not by this framework (zoo/src/test/resources and pyzoo/test/zoo/resources; each test names the reference test that reads
the same file).  Skipped where the reference tree is not on the machine."""
import os

import numpy as np
import pytest
import torch

S = "/root/reference/zoo/src/test/resources/"
P = "/root/reference/pyzoo/test/zoo/resources/"
pytestmark = pytest.mark.skipif(not (os.path.isdir(S) and os.path.isdir(P)), reason="reference test resources are not on this machine")


def test_news20_text_pipeline_and_glove():
    """TextSetSpec.scala / test_text_set.py: read category folders, tokenize -> normalize -> word2idx -> shape_sequence ->
    generate_sample; WordEmbedding over the GloVe excerpt."""
    from zoo.feature.text import TextSet
    from zoo.pipeline.api.keras.layers import WordEmbedding

    ts = TextSet.read(S + "news20")
    assert len(ts.get_texts()) == 5 and sorted(set(ts.get_labels())) == [0, 1, 2]
    out = ts.tokenize().normalize().word2idx(remove_topN=1, max_words_num=50).shape_sequence(len=10).generate_sample()
    wi = out.get_word_index()
    assert len(wi) == 50 and min(wi.values()) == 1
    assert all(s.feature.shape == (10,) and 0 <= s.feature.min() and s.feature.max() <= 50 for s in out.get_samples())
    glove = S + "glove.6B/glove.6B.50d.txt"
    index = WordEmbedding.get_word_index(glove)
    assert len(index) == 21 and "the" in index
    emb = WordEmbedding(glove, {"the": 1, "of": 2, "not-in-the-file": 3}, input_length=4)
    vec = emb(torch.tensor([[1, 2, 3, 0]]))
    assert vec.shape == (1, 4, 50)
    first_line = open(glove).readline().split()
    assert first_line[0] == "the" and np.allclose(vec[0, 0].detach().numpy(), np.asarray(first_line[1:], np.float32), atol=1e-6)
    assert float(vec[0, 2].abs().sum()) == 0.0 and float(vec[0, 3].abs().sum()) == 0.0        # unknown word and padding: zero rows


def test_qa_relations_and_corpus_in_three_formats():
    """RelationsSpec.scala / TextSetSpec.scala: the same four relations as csv, txt and parquet; the corpus as csv and parquet."""
    from zoo.feature.common import Relations
    from zoo.feature.text import TextSet

    rels = [Relations.read(S + "qa/relations.csv"), Relations.read(S + "qa/relations.txt"), Relations.read_parquet(S + "qa/relations.parquet")]
    as_rows = lambda rs: [(r.id1, r.id2, int(r.label)) for r in rs]
    assert as_rows(rels[0]) == [("Q1", "A1", 1), ("Q1", "A2", 0), ("Q2", "A1", 0), ("Q2", "A2", 1)]
    assert as_rows(rels[1]) == as_rows(rels[0]) == as_rows(rels[2])
    q_csv, q_pq = TextSet.read_csv(S + "qa/question_corpus.csv"), TextSet.read_parquet(S + "qa/question_corpus.parquet")
    assert sorted(q_csv.get_texts()) == sorted(q_pq.get_texts()) == ["how old are you?", "what is your project?"]
    assert sorted(q_csv.get_uris()) == ["Q1", "Q2"]
    q = q_csv.tokenize().normalize().word2idx().shape_sequence(len=5)
    import tempfile

    tmp = tempfile.mkdtemp()
    with open(os.path.join(tmp, "answers.csv"), "w") as f:
        f.write("A1,a deep learning framework\nA2,twenty years\n")
    a = TextSet.read_csv(os.path.join(tmp, "answers.csv")).tokenize().normalize().word2idx(existing_map=q.get_word_index()).shape_sequence(len=6)
    pairs = TextSet.from_relation_pairs(rels[0], q, a)
    lists = TextSet.from_relation_lists(rels[0], q, a)
    assert len(pairs.get_samples()) == 2 and len(lists.get_samples()) == 2          # one (positive, negative) pair / one list per question
    assert pairs.get_samples()[0].feature.shape == (2, 11) and lists.get_samples()[0].feature.shape == (2, 11)


def test_image_folders_and_nnimagereader():
    """ImageSetSpec.scala (imagenet folder with class sub-folders) and NNImageReaderSpec.scala / test_nn_image_reader.py (cat_dog)."""
    from zoo.feature.image import ImageSet
    from zoo.pipeline.nnframes import NNImageReader

    s = ImageSet.read(S + "imagenet", with_label=True)
    assert len(s.get_image_features()) == 11 and s.label_map == {"n02110063": 1, "n04370456": 2, "n15075141": 3, "n99999999": 4}
    assert all(f["mat"].ndim == 3 and f["mat"].shape[2] == 3 for f in s.get_image_features())
    resized = ImageSet.read(S + "imagenet", resize_height=32, resize_width=48).get_image_features()
    assert all(f["mat"].shape == (32, 48, 3) for f in resized)
    df = NNImageReader.readImages(P + "cat_dog")
    assert len(df) == 12
    row = df["image"].iloc[0]
    assert set(row) >= {"origin", "height", "width", "nChannels", "mode", "data"} and row["nChannels"] == 3
    gray = ImageSet.read(P + "gray").get_image_features()
    assert len(gray) == 1 and gray[0]["mat"].shape[2] == 3                                 # a gray BMP comes out as 3 channels (IMREAD_COLOR)


def test_pascal_voc_devkit():
    """PascalVoc over the two-image VOCdevkit excerpt (objectdetection dataset specs): 1-based pixel boxes, class ids, difficult flag."""
    from zoo.models.image.objectdetection.dataset import PascalVoc

    voc = PascalVoc("2007", "testcode", S + "VOCdevkit")
    assert voc.image_index() == ["000025", "000055"]
    db = voc.roidb()
    assert db[0]["roi"].shape == (11, 6) and db[1]["roi"].shape == (4, 6) and os.path.exists(db[0]["path"])
    first = db[0]["roi"][0]
    assert voc.classes[int(first[0])] == "cow" and first[1:5].tolist() == [1.0, 83.0, 58.0, 247.0] and first[5] == 0.0


def test_mnist_tfrecords_written_by_tensorflow():
    """test_tf_dataset.py `test_tfdataset_with_tfrecord`: slim-style MNIST records (image/encoded PNG, image/class/label, ...)."""
    from zoo.tfpark import TFDataset
    from zoo.tfpark.tfrecord import read_examples

    exs = list(read_examples(P + "tfrecord/mnist_train.tfrecord"))
    assert len(exs) == 20 and set(exs[0]) == {"image/encoded", "image/format", "image/class/label", "image/height", "image/width"}
    assert exs[0]["image/format"] == [b"png"] and int(exs[0]["image/height"][0]) == 28
    ds = TFDataset.from_tfrecord_file(P + "tfrecord/mnist_train.tfrecord", batch_size=4, features="image/encoded", labels="image/class/label",
                                      validation_file_path=P + "tfrecord/mnist_test.tfrecord")
    batch = next(iter(ds.get_training_data().batches(4, train=False)))
    assert batch.inputs.shape == (4, 28, 28) and batch.inputs.dtype == torch.uint8 and batch.targets.flatten().tolist() == [5, 0, 4, 1]
    assert ds.get_validation_data() is not None
    every = TFDataset.from_tfrecord_file(P + "tfrecord/mnist_train.tfrecord", batch_size=4, labels="image/class/label")
    b = next(iter(every.get_training_data().batches(4, train=False)))
    assert [tuple(t.shape) for t in b.inputs] == [(4, 28, 28), (4, 1), (4, 1)]            # the format tag (a string) is left out


def test_xshard_csv_and_json_folders():
    """test_ray_pandas.py: two csv and two JSON-lines files in one folder -> two shards each."""
    from zoo.xshard.pandas import read_csv, read_json

    csv = read_csv(P + "xshard").collect()
    assert len(csv) == 2 and "location" in csv[0].columns and sum(len(d) for d in csv) == 14
    js = read_json(P + "xshard").collect()
    assert len(js) == 2 and "value" in js[0].columns


def test_recommender_parquet_trains_ncf():
    """RecommenderSpec.scala reads recommender/data.parquet (ML-1M excerpt joined with user and item columns)."""
    import pandas as pd

    from zoo.models.recommendation import NeuralCF

    df = pd.read_parquet(S + "recommender/data.parquet")
    assert df.shape == (458, 8) and {"userId", "itemId", "label", "gender", "age", "occupation", "genres"} <= set(df.columns)
    torch.manual_seed(0)
    ncf = NeuralCF(int(df["userId"].max()), int(df["itemId"].max()), 5, user_embed=8, item_embed=8, hidden_layers=[16, 8], mf_embed=4)
    ncf.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    x, y = df[["userId", "itemId"]].values.astype(np.int64), df["label"].values.astype(np.int64) - 1
    before = ncf.evaluate(x, y, batch_size=128)[0]
    ncf.fit(x, y, batch_size=64, nb_epoch=8)
    assert ncf.evaluate(x, y, batch_size=128)[0] < before


def test_roi_sequence_files_from_the_voc_devkit(tmp_path):
    """RoiImageSeqGenerator -> ROI SequenceFiles -> SSDDataSet chains (RoiByteImageToSeq.scala, SSDDataSet.scala:31-66) over the
    VOCdevkit excerpt: records carry the encoded image + class / box tensors and read back to the annotations."""
    from zoo.models.image.objectdetection import PascalVoc, RoiImageSeqGenerator, SSDDataSet, load_roi_seq_files
    from zoo.models.image.objectdetection.dataset import decode_roi_record, encode_roi_record

    img, roi = decode_roi_record(encode_roi_record(b"\xff\xd8jpegbytes", [[3, 1, 2, 30, 40, 1], [7, 5, 6, 50, 60, 0]]))
    assert img == b"\xff\xd8jpegbytes" and roi.tolist() == [[3, 1, 2, 30, 40, 1], [7, 5, 6, 50, 60, 0]]
    assert decode_roi_record(encode_roi_record(b"x")) == (b"x", None)                          # unlabelled record
    files = RoiImageSeqGenerator.generate(S + "VOCdevkit", str(tmp_path), "voc_2007_testcode", block_size=1, parallel=2)
    assert len(files) == 2 and all(f.endswith(".seq") for f in files)
    feats = load_roi_seq_files(str(tmp_path))
    want = {d["path"]: d["roi"] for d in PascalVoc("2007", "testcode", S + "VOCdevkit").roidb()}
    assert len(feats) == 2 and all(np.array_equal(want[f["uri"]], f["roi"]) and f["bytes"][:2] == b"\xff\xd8" for f in feats)
    np.random.seed(0)
    x, t = next(iter(SSDDataSet.load_ssd_val_set(str(tmp_path), resolution=300, batch_size=2)()))
    assert x.shape == (2, 3, 300, 300) and t.shape == (15, 7) and float(t[:, 2:6].min()) >= 0 and float(t[:, 2:6].max()) <= 1
    assert sorted(t[:, 0].tolist()) == [0.0] * 11 + [1.0] * 4
    x, t = next(iter(SSDDataSet.load_ssd_train_set(str(tmp_path), resolution=300, batch_size=2)()))
    assert x.shape == (2, 3, 300, 300) and t.shape[1] == 7 and len(t) >= 1                        # the random sampler keeps at least one box
    unl = RoiImageSeqGenerator.generate(S + "VOCdevkit/VOC2007/JPEGImages", str(tmp_path / "plain"), None, block_size=10)
    assert len(unl) == 1 and all("roi" not in f for f in load_roi_seq_files(str(tmp_path / "plain")))
