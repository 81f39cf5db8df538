"""Recommendation inference through the POJO API (reference: apps/model-inference-examples/recommendation-inference:
NueralCFJModel.java `extends AbstractInferenceModel` with `preProcess(List<UserItemPair>)`, SimpleDriver.java loads the saved NCF
model and scores user-item pairs).  The NCF model is trained here on synthetic ratings first (the reference loads ncf.bigdl)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", ".."))
import numpy as np
import torch

from zoo.models.recommendation import NeuralCF
from zoo.pipeline.inference import AbstractInferenceModel, JTensor


class UserItemPair:
    def __init__(self, user_id, item_id):
        self.user_id, self.item_id = int(user_id), int(item_id)

    def __repr__(self):
        return "UserItemPair(%d, %d)" % (self.user_id, self.item_id)


class NueralCFJModel(AbstractInferenceModel):
    def pre_process(self, user_item_pairs):
        """pairs -> a list of records, one JTensor [2] = (userId, itemId) each."""
        return [[JTensor.from_ndarray(np.asarray([p.user_id, p.item_id], dtype=np.float32))] for p in user_item_pairs]

    preProcess = pre_process


if __name__ == "__main__":
    torch.manual_seed(0)
    rng = np.random.RandomState(0)
    users, items = 50, 40
    u, i = rng.randint(1, users + 1, 4000), rng.randint(1, items + 1, 4000)
    label = ((u + i) % 5).astype(np.int64)                       # rating class 0..4
    ncf = NeuralCF(users, items, 5, user_embed=16, item_embed=16, hidden_layers=[32, 16], mf_embed=8)
    ncf.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    ncf.fit(np.stack([u, i], 1), label, batch_size=200, nb_epoch=int(os.environ.get("EPOCHS", 25)))
    path = os.path.join(tempfile.mkdtemp(prefix="zoo_ncf_"), "ncf.model")
    ncf.save_model(path)

    model = NueralCFJModel()
    model.load(path)                                             # rcm.load(modelPath) of SimpleDriver.java
    pairs = [UserItemPair(k, k + 3) for k in range(1, 11)]
    results = model.predict(model.pre_process(pairs))
    hits = 0
    for p, r in zip(pairs, results):
        probs = r[0].to_ndarray().reshape(-1)
        hits += int(probs.argmax() == (p.user_id + p.item_id) % 5)
        print(p, "-> class %d  probability %.3f" % (int(probs.argmax()) + 1, float(probs.max())))
    print("pairs whose predicted class is the generating one: %d/%d" % (hits, len(pairs)))
