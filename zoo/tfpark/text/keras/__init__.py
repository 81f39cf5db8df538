"""Text models with a Keras-style surface: NER (BiLSTM-CRF), SequenceTagger (POS + chunk), IntentEntity (joint intent and slot).

Reference: pyzoo/zoo/tfpark/text/keras/{text_model.py:22-70, ner.py:23-70, pos_tagging.py:23-80, intent_extraction.py:23-80},
which wrap the NLP-Architect tf.keras models NERCRF / SequenceTagger / MultiTaskIntentModel in tfpark.KerasModel.  Same
constructor arguments, inputs and outputs; the networks are torch modules and the CRF is a batched linear-chain CRF
(forward algorithm for the likelihood, Viterbi for decoding) that runs on the GPU.

Inputs: words [B, L] int, word_chars [B, L, word_length] int (NER / IntentEntity; optional for SequenceTagger).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


class CRF(nn.Module):
    """Linear-chain CRF over emission scores [B, L, T] with an optional length mask."""

    def __init__(self, num_tags):
        super().__init__()
        self.num_tags = num_tags
        self.trans = nn.Parameter(torch.zeros(num_tags, num_tags))      # trans[i, j]: score of i -> j
        self.start = nn.Parameter(torch.zeros(num_tags))
        self.end = nn.Parameter(torch.zeros(num_tags))
        nn.init.uniform_(self.trans, -0.1, 0.1)

    def _mask(self, em, lengths):
        B, L, _ = em.shape
        if lengths is None:
            return torch.ones(B, L, dtype=torch.bool, device=em.device)
        return torch.arange(L, device=em.device)[None, :] < lengths.reshape(-1, 1).to(em.device)

    def nll(self, em, tags, lengths=None):
        em, mask = em.float(), self._mask(em, lengths)
        B, L, T = em.shape
        tags = tags.long()
        # score of the given path
        first = self.start[tags[:, 0]] + em[:, 0].gather(1, tags[:, :1]).squeeze(1)
        step = self.trans[tags[:, :-1], tags[:, 1:]] + em[:, 1:].gather(2, tags[:, 1:, None]).squeeze(2)
        path = first + (step * mask[:, 1:]).sum(1)
        last = tags.gather(1, (mask.sum(1) - 1).clamp(min=0)[:, None]).squeeze(1)
        path = path + self.end[last]
        # log partition (forward algorithm)
        alpha = self.start[None] + em[:, 0]
        for t in range(1, L):
            nxt = torch.logsumexp(alpha[:, :, None] + self.trans[None], 1) + em[:, t]
            alpha = torch.where(mask[:, t, None], nxt, alpha)
        logz = torch.logsumexp(alpha + self.end[None], 1)
        return (logz - path).mean()

    @torch.no_grad()
    def decode(self, em, lengths=None):
        em, mask = em.float(), self._mask(em, lengths)
        B, L, T = em.shape
        score = self.start[None] + em[:, 0]
        back = []
        for t in range(1, L):
            cand = score[:, :, None] + self.trans[None]
            best, idx = cand.max(1)
            nxt = best + em[:, t]
            score = torch.where(mask[:, t, None], nxt, score)
            back.append(torch.where(mask[:, t, None], idx, torch.arange(T, device=em.device)[None].expand(B, T)))
        tag = (score + self.end[None]).argmax(1)
        out = [tag]
        for idx in reversed(back):
            tag = idx.gather(1, tag[:, None]).squeeze(1)
            out.append(tag)
        return torch.stack(out[::-1], 1)


class _CharWord(nn.Module):
    """word embedding ++ per-word BiLSTM over character embeddings."""

    def __init__(self, word_vocab, char_vocab, word_dim, char_dim, char_lstm_dim, dropout):
        super().__init__()
        self.word = nn.Embedding(word_vocab, word_dim)
        self.char = nn.Embedding(char_vocab, char_dim) if char_vocab else None
        self.char_lstm = nn.LSTM(char_dim, char_lstm_dim, batch_first=True, bidirectional=True) if char_vocab else None
        self.drop = nn.Dropout(dropout)
        self.out_dim = word_dim + (2 * char_lstm_dim if char_vocab else 0)

    def forward(self, words, chars=None):
        w = self.word(words.long())
        if self.char is not None and chars is not None:
            B, L, C = chars.shape
            c = self.char(chars.long().reshape(B * L, C))
            _, (h, _) = self.char_lstm(c)
            w = torch.cat([w, torch.cat([h[0], h[1]], -1).reshape(B, L, -1)], -1)
        return self.drop(w)


class _NerNet(nn.Module):
    def __init__(self, num_entities, word_vocab_size, char_vocab_size, word_length, word_emb_dim, char_emb_dim, tagger_lstm_dim, dropout, crf_mode):
        super().__init__()
        if crf_mode not in ("reg", "pad"):
            raise ValueError("crf_mode should be either 'reg' or 'pad'")
        self.crf_mode = crf_mode
        self.emb = _CharWord(word_vocab_size, char_vocab_size, word_emb_dim, char_emb_dim, char_emb_dim, dropout)
        self.lstm = nn.LSTM(self.emb.out_dim, tagger_lstm_dim, num_layers=2, batch_first=True, bidirectional=True, dropout=dropout)
        self.drop = nn.Dropout(dropout)
        self.out = nn.Linear(2 * tagger_lstm_dim, num_entities)
        self.crf = CRF(num_entities)

    def emissions(self, words, chars, lengths=None):
        h, _ = self.lstm(self.emb(words, chars))
        return self.out(self.drop(h))

    def loss(self, xs, y):
        words, chars = xs[0], xs[1]
        lengths = xs[2].reshape(-1) if (self.crf_mode == "pad" and len(xs) > 2) else None
        return self.crf.nll(self.emissions(words, chars), _ids(y), lengths)

    def predict(self, xs):
        lengths = xs[2].reshape(-1) if (self.crf_mode == "pad" and len(xs) > 2) else None
        tags = self.crf.decode(self.emissions(xs[0], xs[1]), lengths)
        return F.one_hot(tags, self.crf.num_tags).float()        # the reference returns per-token class scores


class _TaggerNet(nn.Module):
    def __init__(self, num_pos_labels, num_chunk_labels, word_vocab_size, char_vocab_size, word_length, feature_size, dropout, classifier):
        super().__init__()
        self.emb = _CharWord(word_vocab_size, char_vocab_size, feature_size, feature_size // 4 if char_vocab_size else 0,
                             feature_size // 4 if char_vocab_size else 0, dropout)
        self.l1 = nn.LSTM(self.emb.out_dim, feature_size, batch_first=True, bidirectional=True)
        self.l2 = nn.LSTM(2 * feature_size, feature_size, batch_first=True, bidirectional=True)
        self.l3 = nn.LSTM(2 * feature_size, feature_size, batch_first=True, bidirectional=True)
        self.drop = nn.Dropout(dropout)
        self.pos = nn.Linear(2 * feature_size, num_pos_labels)
        self.chunk = nn.Linear(2 * feature_size, num_chunk_labels)
        self.crf = CRF(num_chunk_labels) if classifier == "crf" else None

    def _features(self, xs):
        words, chars = xs[0], (xs[1] if len(xs) > 1 else None)
        h1, _ = self.l1(self.emb(words, chars))
        h2, _ = self.l2(self.drop(h1))
        h3, _ = self.l3(self.drop(h2))
        return self.pos(self.drop(h1)), self.chunk(self.drop(h3))

    def loss(self, xs, y):
        pos, chunk = self._features(xs)
        yp, yc = _ids(y[0]), _ids(y[1])
        l = F.cross_entropy(pos.reshape(-1, pos.shape[-1]), yp.reshape(-1))
        if self.crf is not None:
            return l + self.crf.nll(chunk, yc)
        return l + F.cross_entropy(chunk.reshape(-1, chunk.shape[-1]), yc.reshape(-1))

    def predict(self, xs):
        pos, chunk = self._features(xs)
        c = F.one_hot(self.crf.decode(chunk), chunk.shape[-1]).float() if self.crf is not None else torch.softmax(chunk, -1)
        return [torch.softmax(pos, -1), c]


class _IntentNet(nn.Module):
    def __init__(self, num_intents, num_entities, word_vocab_size, char_vocab_size, word_length, word_emb_dim, char_emb_dim, char_lstm_dim,
                 tagger_lstm_dim, dropout):
        super().__init__()
        self.emb = _CharWord(word_vocab_size, char_vocab_size, word_emb_dim, char_emb_dim, char_lstm_dim, dropout)
        self.shared = nn.LSTM(self.emb.out_dim, tagger_lstm_dim, batch_first=True, bidirectional=True)
        self.tagger = nn.LSTM(2 * tagger_lstm_dim, tagger_lstm_dim, batch_first=True, bidirectional=True)
        self.drop = nn.Dropout(dropout)
        self.intent = nn.Linear(2 * tagger_lstm_dim, num_intents)
        self.entity = nn.Linear(2 * tagger_lstm_dim, num_entities)
        self.crf = CRF(num_entities)

    def _features(self, xs):
        h, (hn, _) = self.shared(self.emb(xs[0], xs[1]))
        intent = self.intent(self.drop(torch.cat([hn[0], hn[1]], -1)))
        t, _ = self.tagger(self.drop(h))
        return intent, self.entity(self.drop(t))

    def loss(self, xs, y):
        intent, ent = self._features(xs)
        return F.cross_entropy(intent, _ids(y[0]).reshape(-1)) + self.crf.nll(ent, _ids(y[1]))

    def predict(self, xs):
        intent, ent = self._features(xs)
        return [torch.softmax(intent, -1), F.one_hot(self.crf.decode(ent), ent.shape[-1]).float()]


def _ids(y):
    """Labels as class ids: accepts ids [..] or one-hot [.., C] (the reference's models are fed one-hot tags)."""
    y = torch.as_tensor(y)
    if y.dtype.is_floating_point and y.dim() >= 2 and y.shape[-1] > 1:
        return y.argmax(-1)
    return y.long().squeeze(-1) if (y.dim() >= 2 and y.shape[-1] == 1) else y.long()


class TextKerasModel:
    """fit / evaluate / predict / save_model / load_model over one of the text networks (text_model.py:22-70)."""

    def __init__(self, net, optimizer=None, config=None, device=None):
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.net = net.to(self.device)
        self.config = config or {}
        if optimizer is None:
            self.optimizer = torch.optim.Adam(self.net.parameters(), lr=1e-3)
        elif isinstance(optimizer, torch.optim.Optimizer):
            self.optimizer = optimizer
        else:
            self.optimizer = optimizer(self.net.parameters())

    @staticmethod
    def _as_list(x):
        return [torch.as_tensor(np.asarray(t)) for t in (x if isinstance(x, (list, tuple)) else [x])]

    def _batches(self, xs, ys, batch_size, shuffle):
        n = xs[0].shape[0]
        idx = np.random.permutation(n) if shuffle else np.arange(n)
        for s in range(0, n, batch_size):
            sel = torch.as_tensor(idx[s:s + batch_size])
            yield [t[sel].to(self.device) for t in xs], (None if ys is None else [t[sel].to(self.device) for t in ys])

    def fit(self, x, y, batch_size=32, epochs=1, validation_data=None, distributed=False, **kw):
        xs, ys = self._as_list(x), self._as_list(y)
        history = []
        for _ in range(epochs):
            self.net.train()
            tot, n = 0.0, 0
            for bx, by in self._batches(xs, ys, batch_size, True):
                loss = self.net.loss(bx, by if len(by) > 1 else by[0])
                self.optimizer.zero_grad(set_to_none=True)
                loss.backward()
                nn.utils.clip_grad_norm_(self.net.parameters(), 5.0)
                self.optimizer.step()
                tot, n = tot + float(loss.detach()), n + 1
            rec = {"loss": tot / max(n, 1)}
            if validation_data is not None:
                rec["val_loss"] = self.evaluate(validation_data[0], validation_data[1], batch_size)
            history.append(rec)
        self.history = history
        return self

    def evaluate(self, x, y, batch_per_thread=32, distributed=False, **kw):
        xs, ys = self._as_list(x), self._as_list(y)
        self.net.eval()
        tot, n = 0.0, 0
        with torch.no_grad():
            for bx, by in self._batches(xs, ys, batch_per_thread, False):
                tot, n = tot + float(self.net.loss(bx, by if len(by) > 1 else by[0])) * bx[0].shape[0], n + bx[0].shape[0]
        return tot / max(n, 1)

    def predict(self, x, batch_per_thread=32, distributed=False, **kw):
        xs = self._as_list(x)
        self.net.eval()
        outs = []
        with torch.no_grad():
            for bx, _ in self._batches(xs, None, batch_per_thread, False):
                outs.append(self.net.predict(bx))
        if isinstance(outs[0], list):
            return [torch.cat([o[i] for o in outs]).cpu().numpy() for i in range(len(outs[0]))]
        return torch.cat(outs).cpu().numpy()

    def save_model(self, path):
        torch.save({"class": type(self).__name__, "config": self.config, "state": self.net.state_dict()}, path)

    @classmethod
    def load_model(cls, path):
        ck = torch.load(path, map_location="cpu", weights_only=False)
        if ck["class"] != cls.__name__:
            raise ValueError(f"{path} holds a {ck['class']}, not a {cls.__name__}")
        m = cls(**ck["config"])
        m.net.load_state_dict(ck["state"])
        return m


class NER(TextKerasModel):
    """`NER(num_entities, word_vocab_size, char_vocab_size, word_length=12, word_emb_dim=100, char_emb_dim=30, tagger_lstm_dim=100,
    dropout=0.5, crf_mode='reg', optimizer=None)`; input [words, word_chars] (+ sequence lengths in 'pad' mode)."""

    def __init__(self, num_entities, word_vocab_size, char_vocab_size, word_length=12, word_emb_dim=100, char_emb_dim=30, tagger_lstm_dim=100,
                 dropout=0.5, crf_mode="reg", optimizer=None):
        cfg = dict(num_entities=num_entities, word_vocab_size=word_vocab_size, char_vocab_size=char_vocab_size, word_length=word_length,
                   word_emb_dim=word_emb_dim, char_emb_dim=char_emb_dim, tagger_lstm_dim=tagger_lstm_dim, dropout=dropout, crf_mode=crf_mode)
        super().__init__(_NerNet(**cfg), optimizer, cfg)


class SequenceTagger(TextKerasModel):
    """`SequenceTagger(num_pos_labels, num_chunk_labels, word_vocab_size, char_vocab_size=None, word_length=12, feature_size=100,
    dropout=0.2, classifier='softmax', optimizer=None)`; outputs [pos tags, chunk tags]."""

    def __init__(self, num_pos_labels, num_chunk_labels, word_vocab_size, char_vocab_size=None, word_length=12, feature_size=100, dropout=0.2,
                 classifier="softmax", optimizer=None):
        classifier = classifier.lower()
        if classifier not in ("softmax", "crf"):
            raise AssertionError("classifier should be either softmax or crf")
        cfg = dict(num_pos_labels=num_pos_labels, num_chunk_labels=num_chunk_labels, word_vocab_size=word_vocab_size,
                   char_vocab_size=char_vocab_size, word_length=word_length, feature_size=feature_size, dropout=dropout, classifier=classifier)
        super().__init__(_TaggerNet(**cfg), optimizer, cfg)


class IntentEntity(TextKerasModel):
    """`IntentEntity(num_intents, num_entities, word_vocab_size, char_vocab_size, word_length=12, word_emb_dim=100, char_emb_dim=30,
    char_lstm_dim=30, tagger_lstm_dim=100, dropout=0.2, optimizer=None)`; outputs [intent, entity tags]."""

    def __init__(self, num_intents, num_entities, word_vocab_size, char_vocab_size, word_length=12, word_emb_dim=100, char_emb_dim=30,
                 char_lstm_dim=30, tagger_lstm_dim=100, dropout=0.2, optimizer=None):
        cfg = dict(num_intents=num_intents, num_entities=num_entities, word_vocab_size=word_vocab_size, char_vocab_size=char_vocab_size,
                   word_length=word_length, word_emb_dim=word_emb_dim, char_emb_dim=char_emb_dim, char_lstm_dim=char_lstm_dim,
                   tagger_lstm_dim=tagger_lstm_dim, dropout=dropout)
        super().__init__(_IntentNet(**cfg), optimizer, cfg)
