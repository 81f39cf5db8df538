"""QA ranking with KNRM (WikiQA-shaped): relation pairs for training (RankHinge), relation lists for NDCG / MAP."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch

from zoo.feature.common import Relation
from zoo.feature.text import TextSet
from zoo.feature.text.text_feature import TextFeature
from zoo.models.textmatching import KNRM
from zoo.pipeline.api.keras.objectives import RankHinge

rng = np.random.RandomState(0)
vocab = [f"w{i}" for i in range(60)]
q = TextSet([TextFeature(" ".join(rng.choice(vocab[k * 6:(k + 1) * 6], 4)), uri=f"q{k}") for k in range(10)])
docs, rels = [], []
for k in range(10):
    for j in range(4):
        good = j == 0
        words = rng.choice(vocab[k * 6:(k + 1) * 6] if good else vocab, 8)
        docs.append(TextFeature(" ".join(words), uri=f"d{k}_{j}"))
        rels.append(Relation(f"q{k}", f"d{k}_{j}", 1 if good else 0))
q = q.tokenize().word2idx().shape_sequence(4)
d = TextSet(docs).tokenize().word2idx(existing_map=q.get_word_index()).shape_sequence(8)
pairs, lists = TextSet.from_relation_pairs(rels, q, d), TextSet.from_relation_lists(rels, q, d)
knrm = KNRM(4, 8, vocab_size=max(d.get_word_index().values()) + 1, embed_size=16, kernel_num=11)
opt = torch.optim.Adam(knrm.parameters(), lr=1e-2)
loss_fn = RankHinge()
x = torch.from_numpy(np.stack([s.feature for s in pairs.get_samples()])).long()           # [n_pairs, 2, 12]
for epoch in range(30):
    opt.zero_grad()
    scores = knrm(x.reshape(-1, 12)).reshape(-1, 2)
    loss = loss_fn(scores.reshape(-1, 1))
    loss.backward()
    opt.step()
print(f"loss {float(loss):.4f}  NDCG@3 {knrm.evaluate_ndcg(lists, 3):.3f}  MAP {knrm.evaluate_map(lists):.3f}")
