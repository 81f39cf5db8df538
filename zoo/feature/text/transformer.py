"""Text transformers (reference: pyzoo/zoo/feature/text/transformer.py:28-125; Scala Tokenizer / Normalizer /
WordIndexer / SequenceShaper / TextFeatureToSample)."""
from __future__ import annotations

import re

import numpy as np

from ..common import Preprocessing, Sample


class TextTransformer(Preprocessing):
    def __init__(self, bigdl_type="float", *args):
        pass


class Tokenizer(TextTransformer):
    """Split on whitespace (Tokenizer.scala: `text.split("\\\\s+")`)."""

    def apply(self, f):
        f["tokens"] = [t for t in re.split(r"\s+", f["text"]) if t != ""] if f["text"].strip() else []
        return f


class Normalizer(TextTransformer):
    """Lower-case and strip every non a-z character from each token (Normalizer.scala)."""

    def apply(self, f):
        if "tokens" not in f:
            raise ValueError("TextFeature doesn't contain tokens, please tokenize first")
        f["tokens"] = [re.sub(r"[^a-z]", "", t.lower()) for t in f["tokens"]]
        return f


class WordIndexer(TextTransformer):
    """tokens -> indices with `map` (word -> index); unknown and empty tokens are dropped."""

    def __init__(self, map, bigdl_type="float"):
        self.map = dict(map)

    def apply(self, f):
        if "tokens" not in f:
            raise ValueError("TextFeature doesn't contain tokens, please tokenize first")
        f["indexedTokens"] = [float(self.map[t]) for t in f["tokens"] if t in self.map]
        return f


class SequenceShaper(TextTransformer):
    """Truncate (`pre`: keep the last `len`; `post`: keep the first `len`) or pad at the end with `pad_element`."""

    def __init__(self, len, trunc_mode="pre", pad_element=0, bigdl_type="float"):
        if int(len) <= 0:
            raise ValueError("len should be positive")
        if trunc_mode not in ("pre", "post"):
            raise ValueError("Unknown truncation mode")
        self.len, self.trunc_mode, self.pad_element = int(len), trunc_mode, pad_element

    def apply(self, f):
        if "indexedTokens" not in f:
            raise ValueError("TextFeature doesn't contain indexedTokens, please transform from word to index first")
        idx = list(f["indexedTokens"])
        if len(idx) > self.len:
            idx = idx[len(idx) - self.len:] if self.trunc_mode == "pre" else idx[:self.len]
        else:
            idx = idx + [float(self.pad_element)] * (self.len - len(idx))
        f["indexedTokens"] = idx
        return f


class TextFeatureToSample(TextTransformer):
    def apply(self, f):
        if "indexedTokens" not in f:
            raise ValueError("TextFeature doesn't contain indexedTokens, please transform from word to index first")
        x = np.asarray(f["indexedTokens"], dtype=np.float32)
        f["sample"] = Sample(x, np.asarray([f["label"]], dtype=np.float32)) if "label" in f else Sample(x)
        return f
