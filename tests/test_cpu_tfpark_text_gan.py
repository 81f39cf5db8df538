"""Torch-native TFPark pieces: GANEstimator schedule, CRF, NER / SequenceTagger / IntentEntity, BERT estimators."""
import json

import numpy as np
import pytest
import torch
import torch.nn as nn

from zoo.common.triggers import MaxIteration
from zoo.tfpark import GANEstimator, ZooOptimizer
from zoo.tfpark.gan import wasserstein_discriminator_loss, wasserstein_generator_loss
from zoo.tfpark.text import BERTNER, CRF, NER, BERTClassifier, BERTSQuAD, IntentEntity, SequenceTagger, bert_input_fn


def test_gan_estimator_alternates_phases_and_learns_a_shift(tmp_path):
    torch.manual_seed(0)
    gen = nn.Sequential(nn.Linear(2, 16), nn.ReLU(), nn.Linear(16, 2))
    dis = nn.Sequential(nn.Linear(2, 16), nn.ReLU(), nn.Linear(16, 1))

    class D(nn.Module):
        def __init__(self):
            super().__init__()
            self.net = dis

        def forward(self, data, gen_in):
            return self.net(data)

    def input_fn():
        g = torch.Generator().manual_seed(1)
        for _ in range(20):
            yield torch.randn(64, 2, generator=g), torch.randn(64, 2, generator=g) * 0.1 + torch.tensor([3.0, -2.0])

    def d_loss(real, fake):                   # non-saturating GAN loss (stable without weight clipping)
        return nn.functional.softplus(-real).mean() + nn.functional.softplus(fake).mean()

    est = GANEstimator(gen, D(), lambda fake: nn.functional.softplus(-fake).mean(), d_loss,
                       ZooOptimizer(lambda p: torch.optim.Adam(p, lr=5e-3)), lambda p: torch.optim.Adam(p, lr=5e-3),
                       generator_steps=1, discriminator_steps=2, model_dir=str(tmp_path), device="cpu")
    est.train(input_fn, MaxIteration(600))
    assert est.counter == 600
    assert [p for _, p, _ in est.history[:6]] == ["D", "D", "G", "D", "D", "G"]
    out = est.generate(torch.randn(512, 2))
    m = out.mean(0)
    assert float(m[0]) > 1.0 and float(m[1]) < -0.5            # the generator has moved its output towards the data (3, -2)
    again = GANEstimator(nn.Sequential(nn.Linear(2, 16), nn.ReLU(), nn.Linear(16, 2)), D(), wasserstein_generator_loss, wasserstein_discriminator_loss,
                         torch.optim.Adam, torch.optim.Adam, model_dir=str(tmp_path), device="cpu")
    assert again.counter == 600             # restored from model_dir


def test_crf_likelihood_matches_brute_force_and_viterbi_is_optimal():
    torch.manual_seed(0)
    T, L = 3, 4
    crf = CRF(T)
    with torch.no_grad():
        crf.trans.normal_(); crf.start.normal_(); crf.end.normal_()
    em = torch.randn(1, L, T)
    import itertools

    def score(path):
        s = crf.start[path[0]] + em[0, 0, path[0]]
        for t in range(1, L):
            s = s + crf.trans[path[t - 1], path[t]] + em[0, t, path[t]]
        return s + crf.end[path[-1]]

    paths = list(itertools.product(range(T), repeat=L))
    scores = torch.stack([score(p) for p in paths])
    logz = torch.logsumexp(scores, 0)
    tags = torch.tensor([[0, 2, 1, 1]])
    torch.testing.assert_close(crf.nll(em, tags), logz - score((0, 2, 1, 1)))
    assert tuple(crf.decode(em)[0].tolist()) == paths[int(scores.argmax())]
    # a length mask makes the padded steps irrelevant
    em2 = torch.cat([em, torch.randn(1, 2, T) * 50], 1)
    torch.testing.assert_close(crf.nll(em2, torch.tensor([[0, 2, 1, 1, 0, 0]]), torch.tensor([4])), crf.nll(em, tags))
    assert crf.decode(em2, torch.tensor([4]))[0, :4].tolist() == crf.decode(em)[0].tolist()


def _toy_tagging(n=96, L=8, V=30, seed=0):
    rng = np.random.RandomState(seed)
    words = rng.randint(1, V, (n, L))
    chars = rng.randint(1, 20, (n, L, 6))
    tags = (words % 3)                                   # the tag is a function of the word
    return words, chars, tags


def test_ner_learns_and_roundtrips(tmp_path):
    torch.manual_seed(0); np.random.seed(0)
    words, chars, tags = _toy_tagging()
    m = NER(num_entities=3, word_vocab_size=30, char_vocab_size=20, word_length=6, word_emb_dim=16, char_emb_dim=8, tagger_lstm_dim=16, dropout=0.0,
            optimizer=lambda p: torch.optim.Adam(p, lr=2e-2))
    before = m.evaluate([words, chars], np.eye(3)[tags])
    m.fit([words, chars], np.eye(3)[tags], batch_size=32, epochs=12)
    assert m.evaluate([words, chars], np.eye(3)[tags]) < 0.3 * before
    pred = m.predict([words, chars]).argmax(-1)
    assert (pred == tags).mean() > 0.9
    m.save_model(str(tmp_path / "ner.pt"))
    m2 = NER.load_model(str(tmp_path / "ner.pt"))
    np.testing.assert_array_equal(m2.predict([words, chars]).argmax(-1), pred)
    with pytest.raises(ValueError):
        SequenceTagger.load_model(str(tmp_path / "ner.pt"))
    with pytest.raises(ValueError):
        NER(3, 30, 20, crf_mode="bad")


@pytest.mark.parametrize("classifier", ["softmax", "crf"])
def test_sequence_tagger_two_outputs(classifier):
    torch.manual_seed(0); np.random.seed(0)
    words, chars, tags = _toy_tagging(64)
    m = SequenceTagger(num_pos_labels=3, num_chunk_labels=2, word_vocab_size=30, char_vocab_size=None if classifier == "softmax" else 20,
                       word_length=6, feature_size=16, dropout=0.0, classifier=classifier, optimizer=lambda p: torch.optim.Adam(p, lr=2e-2))
    x = [words] if classifier == "softmax" else [words, chars]
    m.fit(x, [tags, tags % 2], batch_size=32, epochs=8)
    pos, chunk = m.predict(x)
    assert pos.shape == (64, 8, 3) and chunk.shape == (64, 8, 2)
    assert (pos.argmax(-1) == tags).mean() > 0.85
    with pytest.raises(AssertionError):
        SequenceTagger(3, 2, 30, classifier="svm")


def test_intent_entity_joint_model():
    torch.manual_seed(0); np.random.seed(0)
    words, chars, tags = _toy_tagging(64)
    intents = words[:, 0] % 2
    m = IntentEntity(num_intents=2, num_entities=3, word_vocab_size=30, char_vocab_size=20, word_length=6, word_emb_dim=16, char_emb_dim=8,
                     char_lstm_dim=8, tagger_lstm_dim=16, dropout=0.0, optimizer=lambda p: torch.optim.Adam(p, lr=2e-2))
    m.fit([words, chars], [intents, tags], batch_size=32, epochs=10)
    pi, pe = m.predict([words, chars])
    assert pi.shape == (64, 2) and pe.shape == (64, 8, 3)
    assert (pe.argmax(-1) == tags).mean() > 0.85 and (pi.argmax(-1) == intents).mean() > 0.8


_CFG = dict(vocab_size=50, hidden_size=32, num_hidden_layers=2, num_attention_heads=4, intermediate_size=64, max_position_embeddings=16,
            hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)


def _bert_records(n=64, L=12, seed=0):
    rng = np.random.RandomState(seed)
    recs = []
    for _ in range(n):
        ids = rng.randint(2, 50, L)
        label = int(ids[1] % 2)
        ids[0] = 1
        recs.append(({"input_ids": ids, "input_mask": np.ones(L, np.int64), "token_type_ids": np.zeros(L, np.int64)}, label))
    return recs


def test_bert_classifier_train_evaluate_predict(tmp_path):
    torch.manual_seed(0)
    cfg_file = tmp_path / "bert_config.json"
    cfg_file.write_text(json.dumps(_CFG))
    recs = _bert_records()
    fn = bert_input_fn(recs, 12, 16)
    est = BERTClassifier(2, str(cfg_file), optimizer=lambda p: torch.optim.Adam(p, lr=2e-3), model_dir=str(tmp_path / "m"), device="cpu")
    before = est.evaluate(fn, ["loss"])["loss"]
    est.train(fn, steps=60)
    res = est.evaluate(fn, ["acc", "loss"])
    assert res["loss"] < before and res["acc"] > 0.8
    probs = est.predict(bert_input_fn([r[0] for r in recs], 12, 16))
    assert len(probs) == 64 and probs[0].shape == (2,) and abs(float(probs[0].sum()) - 1) < 1e-5
    est2 = BERTClassifier(2, _CFG, model_dir=str(tmp_path / "m"), device="cpu")       # restores the checkpoint
    assert est2.global_step == 60 and est2.evaluate(fn, ["acc"])["acc"] == pytest.approx(res["acc"])
    with pytest.raises(ValueError):
        bert_input_fn(recs, 13, 16)
    with pytest.raises(AssertionError):
        bert_input_fn(recs, 12, 16, features={"input_ids", "bogus"})


def test_bert_ner_and_squad_heads():
    torch.manual_seed(0)
    rng = np.random.RandomState(1)
    L = 10
    ner_recs, sq_recs = [], []
    for _ in range(32):
        ids = rng.randint(2, 50, L)
        mask = np.ones(L, np.int64); mask[7:] = 0
        f = {"input_ids": ids, "input_mask": mask, "token_type_ids": np.zeros(L, np.int64)}
        ner_recs.append((f, ids % 3))
        sq_recs.append((f, {"start_positions": int(np.argmax(ids[:7])), "end_positions": int(np.argmin(ids[:7]))}))
    ner = BERTNER(3, _CFG, optimizer=lambda p: torch.optim.Adam(p, lr=3e-3), device="cpu")
    fn = bert_input_fn(ner_recs, L, 8)
    l0 = ner.evaluate(fn, ["loss"])["loss"]
    ner.train(fn, steps=40)
    assert ner.evaluate(fn, ["loss"])["loss"] < l0
    tags = ner.predict(bert_input_fn([r[0] for r in ner_recs], L, 8))
    assert tags[0].shape == (L,)
    sq = BERTSQuAD(_CFG, optimizer=lambda p: torch.optim.Adam(p, lr=3e-3), device="cpu")
    fn = bert_input_fn(sq_recs, L, 8, labels={"start_positions", "end_positions"})
    l0 = sq.evaluate(fn, ["loss"])["loss"]
    sq.train(fn, steps=30)
    assert sq.evaluate(fn, ["loss"])["loss"] < l0
    p = sq.predict(bert_input_fn([r[0] for r in sq_recs], L, 8))
    assert set(p[0]) == {"start_logits", "end_logits"} and p[0]["start_logits"].shape == (L,)


def test_reference_module_names_resolve():
    """Every public name of every module of the reference's python package resolves here (dev/api_parity.py)."""
    import subprocess
    import sys

    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "dev/api_parity.py"], capture_output=True, text=True, cwd=root)
    last = [l for l in out.stdout.splitlines() if "names present" in l][-1]
    assert "(100.0 %)" in last, out.stdout[-2000:]


def test_compat_helpers():
    from zoo.automl.model.MTNet_keras import AttentionRNNWrapper
    from zoo.pipeline.api.utils import remove_batch, toMultiShape
    from zoo.ray.mxnet.mxnet_runner import MXNetRunner
    from zoo.ray.mxnet.utils import create_trainer_config, find_free_port
    from zoo.tfpark.gan.common import GanOptimMethod
    from zoo.tfpark.tfnet import to_sample_rdd
    from zoo.util.engine import compare_version
    from zoo.util.tf import export_tf

    assert toMultiShape((2, 3)) == [[2, 3]] and remove_batch([(None, 4), (None, 5, 6)]) == [[4], [5, 6]]
    assert create_trainer_config(8, seed=3, extra_config={"a": 1}) == {"batch_size": 8, "optimizer": "sgd", "optimizer_params": {"learning_rate": 0.01},
                                                                      "log_interval": 10, "seed": 3, "a": 1}
    assert 1024 < find_free_port() < 65536
    assert compare_version("2.4.3", "2.2") == 1 and compare_version("1.6", "1.6.0") == 0 and compare_version("2.1", "2.2") == -1
    assert [GanOptimMethod(None, None, 2, 1).phase(i) for i in range(4)] == ["D", "D", "G", "D"]
    s = to_sample_rdd(np.zeros((3, 2)), np.arange(3))
    assert len(s) == 3
    with pytest.raises(NotImplementedError):
        export_tf(None, "/tmp/x", [], [])
    a = AttentionRNNWrapper(nn.LSTM(5, 7, batch_first=True))
    a.return_sequences = True
    assert a(torch.randn(2, 6, 5)).shape == (2, 6, 7)
    r = MXNetRunner().setup_distributed({"DMLC_ROLE": "worker"}, create_trainer_config(4), lambda c, e: [(torch.randn(4, 3), torch.randint(0, 2, (4,)))] * 3,
                                        lambda c: nn.Linear(3, 2))
    st = r.train(2)
    assert st["epoch"] == 2 and st["train_loss"] > 0
    assert MXNetRunner().setup_distributed({"DMLC_ROLE": "server"}, {}, None, None).train() == {}
