"""Variational autoencoder with the Keras-style API (reference apps: apps/variational-autoencoder/*.ipynb -- encoder -> [mean,
log_variance] -> GaussianSampler -> decoder; loss = reconstruction (BCE) + KL divergence, written with autograd CustomLoss /
KLDCriterion).  Digits are replaced by 8x8 two-blob images so it trains in seconds."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

import zoo.pipeline.api.autograd as A
from zoo.pipeline.api.keras.engine.topology import Model
from zoo.pipeline.api.keras.layers import Dense, GaussianSampler, Input, merge

rng = np.random.RandomState(0)
n, side, latent = 4000, 8, 2
x = np.zeros((n, side, side), np.float32)
for i in range(n):
    r, c = rng.randint(0, side - 3, 2)
    x[i, r:r + 3, c:c + 3] = 1.0
x = x.reshape(n, -1)
D = side * side

inp = Input(shape=(D,))
h = Dense(64, activation="relu")(inp)
mean, log_var = Dense(latent)(h), Dense(latent)(h)
z = GaussianSampler()([mean, log_var])
decoder_h, decoder_out = Dense(64, activation="relu"), Dense(D, activation="sigmoid")
recon = decoder_out(decoder_h(z))
# one output tensor [reconstruction | mean | log_variance] so that a single custom loss sees all three
vae = Model(inp, merge([recon, mean, log_var], mode="concat"))


def vae_loss(y_true, y_pred):
    rec, mu, lv = y_pred.slice(1, 0, D), y_pred.slice(1, D, latent), y_pred.slice(1, D + latent, latent)
    eps = 1e-7
    bce = -A.sum(y_true * A.log(rec + eps) + (1.0 - y_true) * A.log(1.0 - rec + eps), axis=1)
    kl = -0.5 * A.sum(1.0 + lv - A.square(mu) - A.exp(lv), axis=1)
    return bce + kl


vae.compile(optimizer="adam", loss=A.CustomLoss(vae_loss, y_pred_shape=(D + 2 * latent,), y_true_shape=(D,)))
before = vae.evaluate(x[:500], x[:500], batch_size=250)[0]
vae.fit(x, x, batch_size=100, nb_epoch=int(os.environ.get("EPOCHS", 15)))
after = vae.evaluate(x[:500], x[:500], batch_size=250)[0]
print("negative ELBO per image: before %.2f after %.2f" % (before, after))
# generate: decode points of the latent grid with the trained decoder layers
import torch

with torch.no_grad():
    device = next(decoder_h.parameters()).device                       # the layers live where fit() trained them
    grid = torch.tensor([[a, b] for a in (-1.5, 0.0, 1.5) for b in (-1.5, 0.0, 1.5)], dtype=torch.float32, device=device)
    imgs = decoder_out(decoder_h(grid)).float().cpu().reshape(-1, side, side)
print("decoded 9 latent points; mean 'ink' per image:", [round(float(i.sum()), 1) for i in imgs], "(a 3x3 blob has 9)")
