"""Seq2seq chatbot (reference example: zoo/.../examples/chatbot/Train.scala + Utils.scala: index the query / answer files with a
dictionary, pad to a fixed length with SENTENCESTART / SENTENCEEND, shared word embedding, 3-layer LSTM encoder and decoder, dense
bridge, generator = TimeDistributed(Dense(vocab)) + log-softmax, TimeDistributedMaskCriterion with padding 0, Adam; after each
epoch `model.infer(query, startSign, maxSeqLen, stopSign, buildOutput)` prints the answers to a few queries).

The corpus here is generated (no network): an answer is the query's words mapped through a fixed "reply" dictionary, reversed."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np
import torch
import torch.nn as nn

from zoo.common.dictionary import ZooDictionary
from zoo.models.seq2seq import Bridge, RNNDecoder, RNNEncoder, Seq2seq
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.api.keras.optimizers import Adam

torch.manual_seed(0)
rng = np.random.RandomState(0)
VOCAB, PAD, START, END = 40, 0, 1, 2
SEQ, EMBED, HIDDEN = 6, int(os.environ.get("EMBED", 32)), int(os.environ.get("HIDDEN", 64))
# the example's idx2w / w2idx files become a ZooDictionary (Train.scala:70); it turns token ids back into words for the printout
idx2w = {PAD: "<pad>", START: "SENTENCESTART", END: "SENTENCEEND"}
idx2w.update({i: "word%02d" % i for i in range(3, VOCAB)})
dictionary = ZooDictionary(index2word=idx2w)
reply = rng.permutation(np.arange(3, VOCAB))                    # word w is answered by reply[w - 3]
n = int(os.environ.get("PAIRS", 3000))
queries = rng.randint(3, VOCAB, (n, SEQ))
answers = reply[queries - 3][:, ::-1][:, :4]                    # answer: replies to the last four words, reversed
dec_in = np.concatenate([np.full((n, 1), START), answers], 1)          # SENTENCESTART a1 a2 a3 a4
target = np.concatenate([answers, np.full((n, 1), END)], 1)            # a1 a2 a3 a4 SENTENCEEND

embedding = nn.Embedding(VOCAB, EMBED, padding_idx=PAD)                 # one embedding shared by encoder and decoder
encoder = RNNEncoder.initialize("lstm", 2, HIDDEN, embedding=embedding)
decoder = RNNDecoder.initialize("lstm", 2, HIDDEN, embedding=embedding)
model = Seq2seq(encoder, decoder, [SEQ], [5], bridge=Bridge.initialize("dense", HIDDEN), generator=nn.Sequential(nn.Linear(HIDDEN, VOCAB), nn.LogSoftmax(-1)))
model.compile(optimizer=Adam(lr=4e-3), loss=SparseCategoricalCrossEntropy(log_prob_as_input=True, padding_value=PAD))

to_word = lambda step: step.argmax(-1)                                  # buildOutput: distribution -> next input token
for epoch in range(int(os.environ.get("EPOCHS", 30))):
    model.fit([queries, dec_in], target, batch_size=100, nb_epoch=1)
hits = 0
for q, want in zip(queries[:50], answers[:50]):
    got = model.infer(q, start_sign=np.array([START]), max_seq_len=5, stop_sign=np.array([END]), build_output=to_word).reshape(-1)
    hits += int(len(got) >= 4 and (got[:4] == want).all())
said = model.infer(queries[0], np.array([START]), 5, np.array([END]), to_word).reshape(-1).tolist()
print("Query>", " ".join(dictionary.decode(queries[0])), "\nAnswer>", " ".join(dictionary.decode(said)),
      "(expected", " ".join(dictionary.decode(answers[0].tolist() + [END])), ")")
print("exact answers on 50 training queries: %d/50" % hits)
