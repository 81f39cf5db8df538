"""GAN losses with the TF-GAN names the reference's examples use (pyzoo/zoo/examples/tfpark/gan/gan_train_and_evaluate.py)."""
import torch
import torch.nn.functional as F


def wasserstein_generator_loss(fake_d_outputs):
    return -fake_d_outputs.mean()


def wasserstein_discriminator_loss(real_d_outputs, fake_d_outputs):
    return fake_d_outputs.mean() - real_d_outputs.mean()


def minimax_discriminator_loss(real_d_outputs, fake_d_outputs, label_smoothing=0.25):
    real = F.binary_cross_entropy_with_logits(real_d_outputs, torch.full_like(real_d_outputs, 1.0 - label_smoothing))
    fake = F.binary_cross_entropy_with_logits(fake_d_outputs, torch.zeros_like(fake_d_outputs))
    return real + fake


def minimax_generator_loss(fake_d_outputs):
    return -F.binary_cross_entropy_with_logits(fake_d_outputs, torch.zeros_like(fake_d_outputs))


def modified_generator_loss(fake_d_outputs):
    """Non-saturating generator loss."""
    return F.binary_cross_entropy_with_logits(fake_d_outputs, torch.ones_like(fake_d_outputs))


modified_discriminator_loss = minimax_discriminator_loss


class GanOptimMethod:
    """(reference: gan/common.py:21-60) one OptimMethod that applies the discriminator optimizer for `d_steps` iterations, then
    the generator optimizer for `g_steps`, over disjoint parameter slices.  The torch GANEstimator keeps two optimizers and
    the same phase arithmetic; this class exposes that arithmetic."""

    def __init__(self, d_optim, g_optim, d_steps, g_steps, g_param_size=None):
        self.d_optim, self.g_optim, self.d_steps, self.g_steps, self.g_param_size = d_optim, g_optim, int(d_steps), int(g_steps), g_param_size

    def phase(self, counter):
        return "D" if counter % (self.d_steps + self.g_steps) < self.d_steps else "G"

    def optimizer_for(self, counter):
        return self.d_optim if self.phase(counter) == "D" else self.g_optim
