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
