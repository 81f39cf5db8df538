"""SessionRecommender: GRU over the session's item embeddings (+ MLP over the purchase history) -> next item."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.models.recommendation import SessionRecommender

rng = np.random.RandomState(0)
items, n = 200, 4000
sess = rng.randint(1, items + 1, (n, 5))
hist = rng.randint(1, items + 1, (n, 3))
nxt = (sess[:, -1] % items) + 1                                    # "next item" = successor of the last one
model = SessionRecommender(items, item_embed=32, rnn_hidden_layers=[40, 20], session_length=5, include_history=True, mlp_hidden_layers=[20], history_length=3)
model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["top5accuracy"])
model.fit([sess, hist], nxt - 1, batch_size=200, nb_epoch=3)
print(model.recommend_for_session((sess[:3], hist[:3]), 4, zero_based_label=False))
