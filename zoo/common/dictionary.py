"""`ZooDictionary`: word <-> index vocabulary (reference: zoo/src/main/scala/com/intel/analytics/zoo/common/ZooDictionary.scala:27-140,
over BigDL's `Dictionary`; the chatbot example builds one from its idx2w / w2idx files, Train.scala:70).

Built either from tokenised sentences -- the `vocab_size` most frequent words are kept, the rest form the discard list -- or from
ready-made maps.  An unknown word maps to index `vocab_size` (one past the last kept word); an unknown index maps to a random
discarded word (or a random kept word when nothing was discarded), which is what the reference does when it decodes."""
from __future__ import annotations

import json
import logging
from collections import Counter

import numpy as np

log = logging.getLogger("zoo.dictionary")


class ZooDictionary:
    def __init__(self, dataset=None, vocab_size=None, index2word=None, word2index=None, seed=None):
        self._rng = np.random.RandomState(seed)
        self._word2index, self._index2word, self._discard = {}, {}, []
        if dataset is not None:
            if vocab_size is None or vocab_size <= 0:
                raise ValueError("vocab_size must be a positive number of words to keep")
            counts = Counter(w for sentence in dataset for w in sentence)
            ranked = sorted(counts.items(), key=lambda kv: (-kv[1], kv[0]))      # frequency first, ties in word order (deterministic)
            kept, dropped = ranked[:vocab_size], ranked[vocab_size:]
            for i, (w, _) in enumerate(kept):
                self._word2index[w], self._index2word[i] = i, w
            self._discard = [w for w, _ in dropped]
        elif index2word is not None or word2index is not None:
            if word2index is None:
                word2index = {w: int(i) for i, w in index2word.items()}
            if index2word is None:
                index2word = {int(i): w for w, i in word2index.items()}
            self._word2index = {w: int(i) for w, i in word2index.items()}
            self._index2word = {int(i): w for i, w in index2word.items()}

    # -- the reference's accessors ------------------------------------------------------------------------------------------
    def get_vocab_size(self):
        return len(self._word2index)

    def get_discard_size(self):
        return len(self._discard)

    def vocabulary(self):
        return [self._index2word[i] for i in sorted(self._index2word)]

    def discard_vocab(self):
        return list(self._discard)

    def word2index(self):
        return dict(self._word2index)

    def index2word(self):
        return dict(self._index2word)

    def get_index(self, word):
        return self._word2index.get(word, len(self._word2index))

    def get_word(self, index):
        w = self._index2word.get(int(index))
        if w is not None:
            return w
        if self._discard:
            return self._discard[self._rng.randint(len(self._discard))]
        if not self._index2word:
            raise KeyError(index)
        keys = sorted(self._index2word)
        return self._index2word[keys[self._rng.randint(len(keys))]]

    def add_word(self, word):
        if word in self._word2index:
            return self._word2index[word]
        i = max(self._index2word, default=-1) + 1
        self._word2index[word], self._index2word[i] = i, word
        if word in self._discard:
            self._discard.remove(word)
        return i

    def encode(self, sentence):
        return [self.get_index(w) for w in sentence]

    def decode(self, indices):
        return [self.get_word(i) for i in indices]

    def print(self):
        for w, i in self._word2index.items():
            log.info("%s -> %d", w, i)

    def print_discard(self):
        for w in self._discard:
            log.info("%s", w)

    # -- persistence (Dictionary.save: dictionary.txt + discard.txt; one JSON file here) -----------------------------------
    def save(self, path):
        with open(path, "w", encoding="utf-8") as f:
            json.dump({"word2index": self._word2index, "discard": self._discard}, f, ensure_ascii=False)
        return self

    @classmethod
    def load(cls, path):
        with open(path, encoding="utf-8") as f:
            d = json.load(f)
        out = cls(word2index=d["word2index"])
        out._discard = list(d.get("discard", []))
        return out

    # camelCase names of the Scala API
    getVocabSize, getDiscardSize, discardVocab, getIndex, getWord, addWord, printDiscard = (
        get_vocab_size, get_discard_size, discard_vocab, get_index, get_word, add_word, print_discard)
