"""BERT estimators: classifier, token tagger (NER) and SQuAD span head over a BERT encoder, with the reference's
`train(input_fn, steps) / evaluate(input_fn, eval_methods, steps) / predict(input_fn)` surface.

Reference: pyzoo/zoo/tfpark/text/estimator/bert_base.py:21-150 (bert_model, bert_input_fn, BERTBaseEstimator),
bert_classifier.py:20-100 (pooled output -> dropout 0.1 -> dense -> softmax; cross entropy), bert_ner.py:20-90 (sequence output
-> dropout -> dense(num_entities); loss masked by input_mask; predictions = argmax), bert_squad.py:21-120 (sequence output ->
dense(2) -> start / end logits; loss = mean of the two cross entropies).  The reference builds a TF graph from Google's BERT
code and a TF checkpoint; here the encoder is torch (`TransformerBlock`: fused QKV / MLP GEMMs on the dense kernels, SDPA
attention), `bert_config_file` is the usual bert_config.json (or a dict) and `init_checkpoint` a torch state_dict file.
"""
from __future__ import annotations

import json
import os
import tempfile

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ....pipeline.api.keras.layers.self_attention import TransformerBlock, _gelu_erf
from ...unsupported import ZooOptimizer

_DEFAULT_CFG = dict(vocab_size=30522, hidden_size=768, num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072,
                    max_position_embeddings=512, type_vocab_size=2, hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1,
                    initializer_range=0.02)


def load_bert_config(cfg):
    if isinstance(cfg, dict):
        out = dict(_DEFAULT_CFG, **cfg)
    elif cfg is None:
        out = dict(_DEFAULT_CFG)
    else:
        with open(cfg) as fh:
            out = dict(_DEFAULT_CFG, **json.load(fh))
    return out


class BertEncoder(nn.Module):
    """`bert_model(...)`: returns (sequence_output [B, L, H], pooled_output [B, H])."""

    def __init__(self, cfg):
        super().__init__()
        c = self.cfg = load_bert_config(cfg)
        h, r = c["hidden_size"], c["initializer_range"]
        self.word = nn.Embedding(c["vocab_size"], h)
        self.pos = nn.Embedding(c["max_position_embeddings"], h)
        self.tok = nn.Embedding(c["type_vocab_size"], h)
        for t in (self.word, self.pos, self.tok):
            nn.init.normal_(t.weight, 0.0, r)
        self.ln = nn.LayerNorm(h, eps=1e-12)
        self.blocks = nn.ModuleList([TransformerBlock(h, c["num_attention_heads"], c["intermediate_size"], c["hidden_dropout_prob"],
                                                      c["attention_probs_dropout_prob"], r, False, _gelu_erf, 1e-12)
                                     for _ in range(c["num_hidden_layers"])])
        self.pooler = nn.Linear(h, h)
        nn.init.normal_(self.pooler.weight, 0.0, r)

    def forward(self, input_ids, input_mask=None, token_type_ids=None):
        B, L = input_ids.shape
        pos = torch.arange(L, device=input_ids.device).unsqueeze(0)
        tt = token_type_ids if token_type_ids is not None else torch.zeros_like(input_ids)
        x = self.word(input_ids.long()) + self.pos(pos) + self.tok(tt.long())
        x = F.dropout(self.ln(x), self.cfg["hidden_dropout_prob"], self.training)
        mask = None
        if input_mask is not None:
            mask = ((1.0 - input_mask.float()) * -10000.0).view(B, 1, 1, L)
        for b in self.blocks:
            x = b(x, mask)
        return x, torch.tanh(self.pooler(x[:, 0].float()))


class _Head(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.bert = BertEncoder(cfg)

    @staticmethod
    def _dense(i, o):
        d = nn.Linear(i, o)
        nn.init.trunc_normal_(d.weight, std=0.02)
        nn.init.zeros_(d.bias)
        return d


class _ClassifierNet(_Head):
    def __init__(self, cfg, num_classes):
        super().__init__(cfg)
        self.out = self._dense(self.bert.cfg["hidden_size"], num_classes)

    def forward(self, f):
        _, pooled = self.bert(f["input_ids"], f.get("input_mask"), f.get("token_type_ids"))
        return self.out(F.dropout(pooled, 0.1, self.training))

    def loss(self, logits, labels, f):
        return F.cross_entropy(logits.float(), labels.long().reshape(-1))

    def predictions(self, logits, f):
        return torch.softmax(logits.float(), -1)


class _NerNet(_Head):
    def __init__(self, cfg, num_entities):
        super().__init__(cfg)
        self.out = self._dense(self.bert.cfg["hidden_size"], num_entities)

    def forward(self, f):
        seq, _ = self.bert(f["input_ids"], f.get("input_mask"), f.get("token_type_ids"))
        return self.out(F.dropout(seq.float(), 0.1, self.training))

    def loss(self, logits, labels, f):
        per_tok = F.cross_entropy(logits.float().reshape(-1, logits.shape[-1]), labels.long().reshape(-1), reduction="none")
        m = f["input_mask"].float().reshape(-1) if f.get("input_mask") is not None else torch.ones_like(per_tok)
        return (per_tok * m).sum() / m.sum().clamp(min=1.0)

    def predictions(self, logits, f):
        return logits.argmax(-1)


class _SquadNet(_Head):
    def __init__(self, cfg):
        super().__init__(cfg)
        self.out = self._dense(self.bert.cfg["hidden_size"], 2)

    def forward(self, f):
        seq, _ = self.bert(f["input_ids"], f.get("input_mask"), f.get("token_type_ids"))
        return self.out(seq.float())          # [B, L, 2]

    def loss(self, logits, labels, f):
        start, end = logits[..., 0], logits[..., 1]
        return 0.5 * (F.cross_entropy(start, labels["start_positions"].long().reshape(-1)) +
                      F.cross_entropy(end, labels["end_positions"].long().reshape(-1)))

    def predictions(self, logits, f):
        out = {"start_logits": logits[..., 0], "end_logits": logits[..., 1]}
        if "unique_ids" in f:
            out["unique_ids"] = f["unique_ids"]
        return out


class _BertData:
    """What `bert_input_fn(...)(mode)` returns: batches of (feature dict, labels)."""

    def __init__(self, records, feature_names, label_names, batch_size, has_label):
        self.records, self.feature_names, self.label_names = records, feature_names, label_names
        self.batch_size, self.has_label = int(batch_size), has_label

    def __len__(self):
        return len(self.records)

    def batches(self, shuffle=False, rng=None):
        idx = np.arange(len(self.records))
        if shuffle:
            (rng or np.random).shuffle(idx)
        for s in range(0, len(idx), self.batch_size):
            recs = [self.records[i] for i in idx[s:s + self.batch_size]]
            feats = [r[0] if self.has_label else r for r in recs]
            f = {k: torch.as_tensor(np.stack([np.asarray(x[k]) for x in feats])) for k in self.feature_names}
            y = None
            if self.has_label:
                labs = [r[1] for r in recs]
                if isinstance(labs[0], dict):
                    y = {k: torch.as_tensor(np.stack([np.asarray(l[k]) for l in labs])) for k in (self.label_names or labs[0].keys())}
                else:
                    y = torch.as_tensor(np.stack([np.asarray(l) for l in labs]))
            yield f, y


def bert_input_fn(rdd, max_seq_length, batch_size, features=("input_ids", "input_mask", "token_type_ids"), extra_features=None,
                  labels=None, label_size=None):
    """`rdd`: iterable of (feature dict, label or dict of labels) for training / evaluation, of feature dicts for prediction;
    every BERT feature has shape [max_seq_length] (bert_base.py:51-105)."""
    features = set(features)
    if not features.issubset({"input_ids", "input_mask", "token_type_ids"}):
        raise AssertionError("features must be a subset of input_ids, input_mask, token_type_ids")
    names = sorted(features) + (sorted(extra_features) if extra_features else [])
    if extra_features is not None and not isinstance(extra_features, dict):
        raise AssertionError("extra_features should be a dictionary")
    if labels is not None and not isinstance(labels, (list, set)):
        raise ValueError("Wrong labels. labels should be a set of label names if you have multiple labels")
    recs = list(rdd)
    for r in recs[:1]:
        f = r[0] if isinstance(r, (tuple, list)) else r
        for k in features:
            if np.asarray(f[k]).shape != (max_seq_length,):
                raise ValueError(f"feature {k} must have shape [{max_seq_length}]")

    def input_fn(mode):
        has_label = mode in ("train", "eval") and bool(recs) and isinstance(recs[0], (tuple, list))
        return _BertData(recs, names, sorted(labels) if labels and len(set(labels)) > 1 else None, batch_size, has_label)

    return input_fn


class BERTBaseEstimator:
    """The base of the BERT estimators (TFEstimator surface: estimator.py:116-332)."""

    def __init__(self, net, optimizer=None, init_checkpoint=None, model_dir=None, device=None, **params):
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.net = net.to(self.device)
        if init_checkpoint is not None:
            sd = torch.load(init_checkpoint, map_location=self.device)
            target = self.net.bert if not any(k.startswith("bert.") for k in sd) else self.net
            target.load_state_dict(sd, strict=False)
        if isinstance(optimizer, ZooOptimizer):
            optimizer = optimizer.optimizer
        if optimizer is None:
            self.optimizer = None
        elif isinstance(optimizer, torch.optim.Optimizer):
            self.optimizer = optimizer
        else:
            self.optimizer = optimizer(self.net.parameters())
        self.model_dir = model_dir or tempfile.mkdtemp()
        self.params = params
        self.global_step = 0
        ck = os.path.join(self.model_dir, "model.pt")
        if os.path.exists(ck):
            self._load(ck)

    def _load(self, path):
        ck = torch.load(path, map_location=self.device, weights_only=False)
        self.net.load_state_dict(ck["net"])
        self.global_step = ck.get("global_step", 0)

    def _save(self):
        os.makedirs(self.model_dir, exist_ok=True)
        torch.save({"net": self.net.state_dict(), "global_step": self.global_step}, os.path.join(self.model_dir, "model.pt"))

    def _to(self, f, y):
        f = {k: v.to(self.device) for k, v in f.items()}
        if isinstance(y, dict):
            y = {k: v.to(self.device) for k, v in y.items()}
        elif y is not None:
            y = y.to(self.device)
        return f, y

    def train(self, input_fn, steps=None):
        if self.optimizer is None:
            raise ValueError("optimizer is required for training")
        data = input_fn("train")
        self.net.train()
        done, loss = 0, None
        rng = np.random.RandomState(self.global_step)
        while steps is None or done < steps:
            for f, y in data.batches(shuffle=True, rng=rng):
                f, y = self._to(f, y)
                loss = self.net.loss(self.net(f), y, f)
                self.optimizer.zero_grad(set_to_none=True)
                loss.backward()
                self.optimizer.step()
                done += 1
                self.global_step += 1
                if steps is not None and done >= steps:
                    break
            if steps is None:
                break            # one pass when no step count is given
        self.last_loss = None if loss is None else float(loss.detach())
        self._save()
        return self

    def evaluate(self, input_fn, eval_methods, steps=None, checkpoint_path=None):
        if checkpoint_path is not None:
            self._load(checkpoint_path)
        data = input_fn("eval")
        self.net.eval()
        correct = total = 0
        loss_sum, n = 0.0, 0
        with torch.no_grad():
            for i, (f, y) in enumerate(data.batches()):
                if steps is not None and i >= steps:
                    break
                f, y = self._to(f, y)
                logits = self.net(f)
                loss_sum += float(self.net.loss(logits, y, f))
                n += 1
                if not isinstance(y, dict):
                    pred = logits.argmax(-1)
                    if pred.shape == y.shape:
                        m = f["input_mask"].bool() if (pred.dim() == 2 and f.get("input_mask") is not None) else torch.ones_like(pred, dtype=torch.bool)
                        correct += int(((pred == y) & m).sum())
                        total += int(m.sum())
        out = {}
        for name in eval_methods:
            key = name.lower() if isinstance(name, str) else type(name).__name__.lower()
            if key in ("acc", "accuracy", "top1accuracy"):
                out[name if isinstance(name, str) else key] = correct / max(total, 1)
            elif key == "loss":
                out[name if isinstance(name, str) else key] = loss_sum / max(n, 1)
            else:
                raise ValueError(f"unsupported eval method {name!r} (acc, loss)")
        return out

    def predict(self, input_fn, predict_keys=None, checkpoint_path=None):
        if checkpoint_path is not None:
            self._load(checkpoint_path)
        data = input_fn("infer")
        self.net.eval()
        outs = []
        with torch.no_grad():
            for f, _ in data.batches():
                f, _ = self._to(f, None)
                p = self.net.predictions(self.net(f), f)
                if isinstance(p, dict):
                    keys = predict_keys or list(p.keys())
                    for i in range(f["input_ids"].shape[0]):
                        outs.append({k: p[k][i].cpu().numpy() for k in keys})
                else:
                    outs.extend(list(p.cpu().numpy()))
        return outs


class BERTClassifier(BERTBaseEstimator):
    """`BERTClassifier(num_classes, bert_config_file, init_checkpoint=None, use_one_hot_embeddings=False, optimizer=None,
    model_dir=None)`; predictions are class probabilities."""

    def __init__(self, num_classes, bert_config_file, init_checkpoint=None, use_one_hot_embeddings=False, optimizer=None, model_dir=None, **kw):
        super().__init__(_ClassifierNet(bert_config_file, num_classes), optimizer, init_checkpoint, model_dir, num_classes=num_classes, **kw)


class BERTNER(BERTBaseEstimator):
    """Token classification; predictions are entity ids per token."""

    def __init__(self, num_entities, bert_config_file, init_checkpoint=None, use_one_hot_embeddings=False, optimizer=None, model_dir=None, **kw):
        super().__init__(_NerNet(bert_config_file, num_entities), optimizer, init_checkpoint, model_dir, num_entities=num_entities, **kw)


class BERTSQuAD(BERTBaseEstimator):
    """Span extraction; labels are {"start_positions", "end_positions"}, predictions {"start_logits", "end_logits"[, "unique_ids"]}."""

    def __init__(self, bert_config_file, init_checkpoint=None, use_one_hot_embeddings=False, optimizer=None, model_dir=None, **kw):
        super().__init__(_SquadNet(bert_config_file), optimizer, init_checkpoint, model_dir, **kw)
