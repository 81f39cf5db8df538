"""TextSet (reference: pyzoo/zoo/feature/text/text_set.py:23-483, zoo/.../feature/text/TextSet.scala): a collection of
TextFeatures with the pipeline `tokenize -> normalize -> word2idx -> shape_sequence -> generate_sample`, vocabulary
save / load, directory / csv / parquet readers, and the pair / list generators used by the ranking models.
Local and Distributed sets are the same in-memory list here (no Spark)."""
from __future__ import annotations

import os
from collections import Counter

import numpy as np

from ..common import Sample
from .text_feature import TextFeature
from .transformer import Normalizer, SequenceShaper, TextFeatureToSample, Tokenizer, WordIndexer


class TextSet:
    def __init__(self, features=None, word_index=None):
        self.features = list(features or [])
        self.word_index = word_index

    # ---- kinds ---------------------------------------------------------------------------------
    def is_local(self):
        return True

    def is_distributed(self):
        return False

    def to_distributed(self, sc=None, partition_num=4):
        return self

    def to_local(self):
        return self

    # ---- vocabulary ------------------------------------------------------------------------------
    def get_word_index(self):
        return self.word_index

    def set_word_index(self, vocab):
        self.word_index = dict(vocab)
        return self

    def save_word_index(self, path):
        """One `word index` pair per line (TextSet.scala `saveWordIndex`)."""
        if self.word_index is None:
            raise ValueError("wordIndex is null, please call word2idx or setWordIndex first")
        with open(path, "w", encoding="utf-8") as f:
            for w, i in self.word_index.items():
                f.write(f"{w} {i}\n")
        return self

    def load_word_index(self, path):
        wi = {}
        with open(path, encoding="utf-8") as f:
            for line in f:
                w, i = line.rstrip("\n").rsplit(" ", 1)
                wi[w] = int(i)
        self.word_index = wi
        return self

    def generate_word_index_map(self, remove_topN=0, max_words_num=-1, min_freq=1, existing_map=None):
        """Words sorted by descending frequency; drop the `remove_topN` most frequent, keep at most `max_words_num`,
        keep only frequency >= `min_freq`; indices start at 1 (0 is padding); an `existing_map` is extended."""
        cnt = Counter(t for f in self.features for t in (f.get("tokens") or []) if t != "")
        words = [w for w, c in sorted(cnt.items(), key=lambda kv: (-kv[1], kv[0])) if c >= min_freq]
        words = words[remove_topN:]
        if max_words_num > 0:
            words = words[:max_words_num]
        wi = dict(existing_map or {})
        nxt = max(wi.values(), default=0) + 1
        for w in words:
            if w not in wi:
                wi[w] = nxt
                nxt += 1
        self.word_index = wi
        return wi

    # ---- getters ---------------------------------------------------------------------------------
    def get_texts(self):
        return [f.get("text") for f in self.features]

    def get_uris(self):
        return [f.get("uri") for f in self.features]

    def get_labels(self):
        return [f.get("label", -1) for f in self.features]

    def get_predicts(self):
        return [(f.get("uri"), f.get("predict")) for f in self.features]

    def get_samples(self):
        return [f.get("sample") for f in self.features]

    def random_split(self, weights):
        w = np.asarray(weights, dtype=np.float64)
        idx = np.random.permutation(len(self.features))
        cuts = (np.cumsum(w / w.sum()) * len(idx)).round().astype(int)
        out, s = [], 0
        for c in cuts:
            out.append(TextSet([self.features[i] for i in idx[s:c]], self.word_index))
            s = c
        return out

    # ---- pipeline ----------------------------------------------------------------------------------
    def transform(self, transformer):
        return TextSet([transformer(_copy(f)) for f in self.features], self.word_index)

    def tokenize(self):
        return self.transform(Tokenizer())

    def normalize(self):
        return self.transform(Normalizer())

    def word2idx(self, remove_topN=0, max_words_num=-1, min_freq=1, existing_map=None):
        if self.word_index is None or existing_map is not None:
            self.generate_word_index_map(remove_topN, max_words_num, min_freq, existing_map)
        return self.transform(WordIndexer(self.word_index))

    def shape_sequence(self, len, trunc_mode="pre", pad_element=0):
        return self.transform(SequenceShaper(len, trunc_mode, pad_element))

    def generate_sample(self):
        return self.transform(TextFeatureToSample())

    def to_arrays(self):
        """(x [n, len], y [n]) from the generated samples (labels omitted when absent)."""
        xs = np.stack([np.asarray(f["indexedTokens"], dtype=np.int64) for f in self.features])
        if all("label" in f for f in self.features):
            return xs, np.asarray([f["label"] for f in self.features], dtype=np.int64)
        return xs, None

    def to_feature_set(self, **kw):
        from ..common import FeatureSet

        samples = self.get_samples()
        if samples and samples[0] is not None:
            return FeatureSet.sample_rdd(samples, **kw)
        x, y = self.to_arrays()
        return FeatureSet.numpy(x, y, **kw)

    def to_query_batches(self):
        """For `Ranker.evaluate_ndcg/map` over a set built by `from_relation_lists`: (features, labels) per query."""
        return [(np.asarray(s.features[0]), np.asarray(s.labels[0]).reshape(-1)) for s in self.get_samples()]

    # ---- readers -----------------------------------------------------------------------------------
    @classmethod
    def read(cls, path, sc=None, min_partitions=1, bigdl_type="float"):
        """`path/<category>/<file>.txt`: one sub-directory per class; labels are 0-based in sorted category order."""
        feats = []
        cats = sorted(d for d in os.listdir(path) if os.path.isdir(os.path.join(path, d)))
        for lab, c in enumerate(cats):
            for fn in sorted(os.listdir(os.path.join(path, c))):
                fp = os.path.join(path, c, fn)
                if os.path.isfile(fp):
                    with open(fp, encoding="utf-8", errors="ignore") as f:
                        feats.append(TextFeature(f.read(), lab, fp))
        return cls(feats)

    @classmethod
    def read_csv(cls, path, sc=None, min_partitions=1, bigdl_type="float"):
        """Each line `id,text` (no header): uri = id."""
        feats = []
        with open(path, encoding="utf-8") as f:
            for line in f:
                line = line.rstrip("\n")
                if line:
                    i, t = line.split(",", 1)
                    feats.append(TextFeature(t, uri=i))
        return cls(feats)

    @classmethod
    def read_parquet(cls, path, sc=None, bigdl_type="float"):
        import pandas as pd

        df = pd.read_parquet(path)
        return cls([TextFeature(str(r["text"]), uri=str(r["id"])) for _, r in df.iterrows()])

    # ---- relation generators (text_set.py:369-433) -----------------------------------------------------------
    @staticmethod
    def _corpus_map(corpus):
        m = {}
        for f in corpus.features:
            if "indexedTokens" not in f:
                raise ValueError("corpus must have been transformed by word2idx and shape_sequence")
            m[f["uri"]] = np.asarray(f["indexedTokens"], dtype=np.float32)
        return m

    @classmethod
    def from_relation_pairs(cls, relations, corpus1, corpus2, bigdl_type="float"):
        """Pairwise training set: for every (id1, positive id2, negative id2) a sample with feature
        `[2, len1 + len2]` (positive row first) and label `[1, 0]`."""
        from ..common import Relations

        if not isinstance(relations, list):
            raise TypeError("relations should be RDD or list of Relation")
        m1, m2 = cls._corpus_map(corpus1), cls._corpus_map(corpus2)
        feats = []
        for p in Relations.generate_relation_pairs(relations):
            id1, pos, neg = p
            x = np.stack([np.concatenate([m1[id1], m2[pos]]), np.concatenate([m1[id1], m2[neg]])])
            f = TextFeature(uri=f"{id1}{pos}{neg}")
            f["sample"] = Sample(x, np.array([1.0, 0.0], dtype=np.float32))
            feats.append(f)
        return cls(feats)

    @classmethod
    def from_relation_lists(cls, relations, corpus1, corpus2, bigdl_type="float"):
        """Ranking evaluation set: relations grouped by id1; per group a sample with feature `[list_len, len1 + len2]`
        and label `[list_len, 1]`."""
        if not isinstance(relations, list):
            raise TypeError("relations should be RDD or list of Relation")
        m1, m2 = cls._corpus_map(corpus1), cls._corpus_map(corpus2)
        groups = {}
        for r in relations:
            id1, id2, label = r.to_tuple() if hasattr(r, "to_tuple") else r
            groups.setdefault(id1, []).append((id2, label))
        feats = []
        for id1, lst in groups.items():
            x = np.stack([np.concatenate([m1[id1], m2[id2]]) for id2, _ in lst])
            y = np.asarray([[float(l)] for _, l in lst], dtype=np.float32)
            f = TextFeature(uri=str(id1))
            f["sample"] = Sample(x, y)
            feats.append(f)
        return cls(feats)

    def __len__(self):
        return len(self.features)


def _copy(f):
    g = TextFeature()
    g.update(f)
    return g


class LocalTextSet(TextSet):
    def __init__(self, texts=None, labels=None, jvalue=None, bigdl_type="float"):
        feats = []
        for i, t in enumerate(texts or []):
            feats.append(TextFeature(t, None if labels is None else labels[i]))
        super().__init__(feats)


DistributedTextSet = LocalTextSet
