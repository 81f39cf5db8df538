"""TextMatcher base (reference: pyzoo/zoo/models/textmatching/text_matcher.py:24-50): holds the embedding settings and
mixes in the ranking metrics (NDCG / MAP)."""
from __future__ import annotations

from ..common.ranker import Ranker
from ..common.zoo_model import ZooModel


class TextMatcher(ZooModel, Ranker):
    def __init__(self, text1_length, vocab_size, embed_size=300, embed_weights=None, train_embed=True, target_mode="ranking", bigdl_type="float"):
        super().__init__()
        self.text1_length, self.vocab_size, self.embed_size = int(text1_length), int(vocab_size), int(embed_size)
        self.embed_weights, self.train_embed = embed_weights, train_embed
        if target_mode not in ("ranking", "classification"):
            raise ValueError("target_mode should be either ranking or classification, but got " + str(target_mode))
        self.target_mode = target_mode
