"""TextClassifier (reference: pyzoo/zoo/models/textclassification/text_classifier.py:29-180,
TextClassifier.scala:34-70): WordEmbedding (GloVe) -> CNN | LSTM | GRU encoder -> Dense(128) -> Dropout(0.2) -> ReLU
-> Dense(class_num, softmax).  `embedding_file=None` with `vocab_size/embed_dim` builds a trainable embedding instead
(there are no pre-trained vectors on a box without network)."""
from __future__ import annotations

from ...pipeline.api.keras.engine.topology import Sequential
from ...pipeline.api.keras.layers import (GRU, LSTM, Activation, Convolution1D, Dense, Dropout, Embedding, GlobalMaxPooling1D,
                                          WordEmbedding)
from ..common.zoo_model import ZooModel


class TextClassifier(ZooModel):
    def __init__(self, class_num, embedding_file=None, word_index=None, sequence_length=500, encoder="cnn", encoder_output_dim=256,
                 vocab_size=None, embed_dim=None, **kwargs):
        kwargs.pop("bigdl_type", None)
        if "token_length" in kwargs:
            import warnings

            embed_dim = embed_dim or kwargs.pop("token_length")
            warnings.warn('The "token_length" argument in TextClassifier has been deprecated since 0.3.0; pass "embedding_file" and "word_index".')
        if kwargs:
            raise TypeError("Wrong arguments for TextClassifier: " + str(kwargs))
        super().__init__()
        self.class_num, self.sequence_length, self.encoder, self.encoder_output_dim = int(class_num), int(sequence_length), encoder, int(encoder_output_dim)
        self._record_config(class_num=self.class_num, embedding_file=embedding_file, word_index=word_index, sequence_length=self.sequence_length,
                            encoder=encoder, encoder_output_dim=self.encoder_output_dim, vocab_size=vocab_size, embed_dim=embed_dim)
        if embedding_file is not None:
            self.embedding = WordEmbedding(embedding_file, word_index, input_length=self.sequence_length)
        elif vocab_size and embed_dim:
            self.embedding = Embedding(int(vocab_size), int(embed_dim), input_length=self.sequence_length)
        else:
            raise ValueError("TextClassifier needs embedding_file (GloVe text file) or vocab_size + embed_dim")
        self.model = self.build_model()

    def build_model(self):
        model = Sequential()
        model.add(self.embedding)
        enc = self.encoder.lower()
        if enc == "cnn":
            model.add(Convolution1D(self.encoder_output_dim, 5, activation="relu")).add(GlobalMaxPooling1D())
        elif enc == "lstm":
            model.add(LSTM(self.encoder_output_dim))
        elif enc == "gru":
            model.add(GRU(self.encoder_output_dim))
        else:
            raise ValueError("Unsupported encoder for TextClassifier: " + self.encoder)
        model.add(Dense(128)).add(Dropout(0.2)).add(Activation("relu")).add(Dense(self.class_num, activation="softmax"))
        return model

    def forward(self, x):
        return self.model(x)

    def logits_kind(self):
        return self.model.logits_kind()

    def forward_logits(self, x):
        return self.model.forward_logits(x)

    def _xy(self, x):
        if hasattr(x, "to_arrays"):       # a TextSet
            return x.to_arrays()
        return x

    def fit(self, x, y=None, batch_size=32, nb_epoch=10, validation_data=None, **kw):
        x = self._xy(x)
        if isinstance(x, tuple) and y is None:
            x, y = x
        if validation_data is not None:
            validation_data = self._xy(validation_data)
        return super().fit(x, y, batch_size=batch_size, nb_epoch=nb_epoch, validation_data=validation_data)

    def evaluate(self, x, y=None, batch_size=32):
        x = self._xy(x)
        if isinstance(x, tuple) and y is None:
            x, y = x
        return super().evaluate(x, y, batch_size=batch_size)

    def predict(self, x, batch_per_thread=4, **kw):
        x = self._xy(x)
        if isinstance(x, tuple):
            x = x[0]
        return super().predict(x, batch_per_thread=batch_per_thread, **kw)
