"""GANEstimator (reference example: pyzoo/zoo/examples/tensorflow/tfpark/gan/gan_train_and_evaluate.py: MNIST GAN with
TF-GAN losses, generator_steps=1, discriminator_steps=1, checkpoints in model_dir).  Here on a 2-D toy distribution."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import torch
import torch.nn as nn

from zoo.common.triggers import MaxIteration
from zoo.tfpark import GANEstimator, ZooOptimizer
from zoo.tfpark.gan import modified_discriminator_loss, modified_generator_loss

torch.manual_seed(0)
NOISE = 8


class Discriminator(nn.Module):
    def __init__(self):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(2, 32), nn.LeakyReLU(0.2), nn.Linear(32, 32), nn.LeakyReLU(0.2), nn.Linear(32, 1))

    def forward(self, data, generator_inputs):
        return self.net(data)


def input_fn():
    g = torch.Generator().manual_seed(1)
    for _ in range(50):
        angle = torch.rand(128, generator=g) * 6.2832
        real = torch.stack([angle.cos(), angle.sin()], 1) * 2.0 + 0.05 * torch.randn(128, 2, generator=g)   # a ring of radius 2
        yield torch.randn(128, NOISE, generator=g), real


gen = nn.Sequential(nn.Linear(NOISE, 32), nn.ReLU(), nn.Linear(32, 32), nn.ReLU(), nn.Linear(32, 2))
opt = lambda p: torch.optim.Adam(p, lr=2e-3, betas=(0.5, 0.999))
est = GANEstimator(gen, Discriminator(), modified_generator_loss, lambda real, fake: modified_discriminator_loss(real, fake, label_smoothing=0.0),
                   ZooOptimizer(opt), ZooOptimizer(opt), generator_steps=1, discriminator_steps=1, model_dir=tempfile.mkdtemp(), device="cpu")
est.train(input_fn, MaxIteration(int(os.environ.get("STEPS", 1500))))
samples = est.generate(torch.randn(2000, NOISE))
radius = samples.norm(dim=1)
print("generated radius mean %.2f std %.2f (data: 2.00 / 0.05); checkpoint in %s" % (radius.mean(), radius.std(), est.model_dir))
