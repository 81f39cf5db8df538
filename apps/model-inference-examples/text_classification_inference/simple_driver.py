"""SimpleDriver of the reference's example (SimpleDriver.java): load the model, classify a few texts, print class and probability.
Self-contained: writes a small GloVe-format file and trains a tiny CNN text classifier on two synthetic topics first
(the reference trains its model with text-classification-training/TextClassificationTrainer.scala), then also exercises the
web service in-process."""
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, "..", "..", ".."))
import numpy as np
import torch
import torch.nn as nn
from text_classification_model import TextClassificationModel

work = tempfile.mkdtemp(prefix="zoo_textcls_")
rng = np.random.RandomState(0)
topics = {0: ["rocket", "orbit", "launch", "satellite", "moon", "shuttle"], 1: ["engine", "car", "brake", "tire", "dealer", "mileage"]}
common = ["the", "a", "of", "to", "and", "in", "is", "it", "that", "for", "on", "with"]
emb_path = os.path.join(work, "glove.6B.50d.txt")
with open(emb_path, "w") as f:                                   # most frequent first, as in the real file
    for w in common + topics[0] + topics[1]:
        f.write(w + " " + " ".join("%.5f" % v for v in rng.randn(50) * 0.5) + "\n")

model = TextClassificationModel(2, 10, 20, emb_path)


def sentence(label):
    words = list(rng.choice(topics[label], 5)) + list(rng.choice(common, 6))
    return " ".join(rng.permutation(words))


texts, labels = zip(*[(sentence(k % 2), k % 2) for k in range(400)])
x = torch.from_numpy(np.stack([model.preprocess(t).to_ndarray() for t in texts]))
y = torch.tensor(labels)
torch.manual_seed(0)
net = nn.Sequential(nn.Conv1d(20, 32, 5, padding=2), nn.ReLU(), nn.AdaptiveMaxPool1d(1), nn.Flatten(), nn.Linear(32, 2), nn.Softmax(dim=-1))
opt = torch.optim.Adam(net.parameters(), 1e-2)
for _ in range(int(os.environ.get("TRAIN_STEPS", 30))):
    opt.zero_grad()
    loss = nn.functional.nll_loss(torch.log(net(x) + 1e-9), y)
    loss.backward()
    opt.step()
model_path = os.path.join(work, "text-classification.pt")
torch.save(net.eval(), model_path)

model.loadPyTorch(model_path)
for text in ("the shuttle reached orbit after the launch", "the dealer replaced the brake and the tire"):
    out = model.predict([[model.preprocess(text)]])[0][0].to_ndarray().reshape(-1)
    print("class %d  probability %.3f  <- %s" % (int(out.argmax()), float(out.max()), text))

from fastapi.testclient import TestClient  # noqa: E402
from web_service import create_app  # noqa: E402

client = TestClient(create_app(emb_path, model_path, sequence_length=20))
print(client.get("/").text, "|", client.get("/greetings", params={"name": "zoo"}).text)
print(client.post("/predict", content="a satellite and a rocket on the moon").text.strip().replace("\n", "  "))
