from .text_feature import TextFeature  # noqa: F401
from .text_set import DistributedTextSet, LocalTextSet, TextSet  # noqa: F401
from .transformer import Normalizer, SequenceShaper, TextFeatureToSample, TextTransformer, Tokenizer, WordIndexer  # noqa: F401
