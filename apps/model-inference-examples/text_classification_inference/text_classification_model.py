"""`TextClassificationModel extends AbstractInferenceModel` + `TextProcessor` of the reference's web-service example
(apps/model-inference-examples/text-classification-inference: TextClassificationModel.java, TextProcessor.java):
text -> lower-cased tokens -> the `stop_words_count` most frequent words of the embedding file dropped -> the first
`sequence_length` tokens -> their GloVe vectors (zeros for unknown words and padding) -> JTensor [sequence_length, dim]."""
import os
import re
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", ".."))
import numpy as np

from zoo.pipeline.inference import AbstractInferenceModel, JTensor


class TextProcessor:
    def __init__(self, stop_words_count, sequence_length, embedding_file_path):
        self.sequence_length = int(sequence_length)
        self.vectors, order = {}, []
        with open(embedding_file_path, encoding="utf-8") as f:
            for line in f:                                   # GloVe text format: word v1 v2 ...; the file is sorted by frequency
                parts = line.rstrip().split(" ")
                self.vectors[parts[0]] = np.asarray(parts[1:], dtype=np.float32)
                order.append(parts[0])
        self.dim = len(next(iter(self.vectors.values())))
        self.stop_words = set(order[:int(stop_words_count)])

    def tokenize(self, text):
        return [t for t in re.split(r"[^a-z0-9']+", text.lower()) if t]

    def preprocess(self, text):
        tokens = [t for t in self.tokenize(text) if t not in self.stop_words][:self.sequence_length]
        out = np.zeros((self.sequence_length, self.dim), dtype=np.float32)
        for i, t in enumerate(tokens):
            v = self.vectors.get(t)
            if v is not None:
                out[i] = v
        return JTensor.from_ndarray(out)


class TextClassificationModel(AbstractInferenceModel):
    def __init__(self, supported_concurrent_num, stop_words_count, sequence_length, embedding_file_path):
        super().__init__(supported_concurrent_num)
        self.text_processor = TextProcessor(stop_words_count, sequence_length, embedding_file_path)

    def preprocess(self, text):
        return self.text_processor.preprocess(text)
