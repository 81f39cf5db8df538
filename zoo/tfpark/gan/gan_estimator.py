"""GANEstimator: alternating generator / discriminator training with the reference's schedule.

Reference: pyzoo/zoo/tfpark/gan/gan_estimator.py:28-170.  There a TF graph holds both networks and a global step; a step is
a DISCRIMINATOR step while `counter % (discriminator_steps + generator_steps) < discriminator_steps`, else a generator
step; each phase only updates its own variables; checkpoints go to `model_dir`.  Here generator and discriminator are torch
modules (or zero-argument factories of modules), the loss functions keep their signatures
(`generator_loss_fn(fake_d_outputs)`, `discriminator_loss_fn(real_d_outputs, fake_d_outputs)`), optimizers are torch
optimizers, factories `params -> optimizer`, or a `ZooOptimizer` around either, and `input_fn()` yields
(generator_inputs, real_data) batches (a `TFDataset`, a torch `DataLoader`, or any iterable).
"""
from __future__ import annotations

import inspect
import os
import tempfile

import torch
import torch.nn as nn

from ...common.triggers import MaxIteration, ZooTrigger
from ..unsupported import ZooOptimizer


def _module(fn):
    if isinstance(fn, nn.Module):
        return fn
    m = fn()
    if not isinstance(m, nn.Module):
        raise TypeError("generator_fn / discriminator_fn must be a torch module or a zero-argument factory returning one")
    return m


def _optimizer(opt, params):
    if isinstance(opt, ZooOptimizer):
        opt = opt.optimizer
    if isinstance(opt, torch.optim.Optimizer):
        return opt
    if isinstance(opt, type) and issubclass(opt, torch.optim.Optimizer):
        return opt(params)
    if callable(opt):
        return opt(params)
    raise TypeError("optimizer must be a torch optimizer, an optimizer class, or a callable params -> optimizer")


def _call_maybe_with_counter(fn, counter, *args):
    """The reference passes the global step to functions that declare a `counter` argument (gan_estimator.py:65-70)."""
    try:
        wants = "counter" in inspect.signature(fn.forward if isinstance(fn, nn.Module) else fn).parameters
    except (TypeError, ValueError):
        wants = False
    return fn(*args, counter=counter) if wants else fn(*args)


class GANEstimator:
    def __init__(self, generator_fn, discriminator_fn, generator_loss_fn, discriminator_loss_fn, generator_optimizer,
                 discriminator_optimizer, generator_steps=1, discriminator_steps=1, model_dir=None, session_config=None, device=None):
        self.generator, self.discriminator = _module(generator_fn), _module(discriminator_fn)
        self._generator_loss_fn, self._discriminator_loss_fn = generator_loss_fn, discriminator_loss_fn
        self._generator_steps, self._discriminator_steps = int(generator_steps), int(discriminator_steps)
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.generator.to(self.device)
        self.discriminator.to(self.device)
        self._gen_opt = _optimizer(generator_optimizer, self.generator.parameters())
        self._dis_opt = _optimizer(discriminator_optimizer, self.discriminator.parameters())
        self.model_dir = model_dir or tempfile.mkdtemp()
        self.checkpoint_path = os.path.join(self.model_dir, "model")
        self.counter = 0
        self.history = []          # (step, phase, loss)
        if os.path.exists(self.checkpoint_path + ".pt"):
            self._restore()

    def _restore(self):
        ck = torch.load(self.checkpoint_path + ".pt", map_location=self.device, weights_only=False)
        self.generator.load_state_dict(ck["generator"])
        self.discriminator.load_state_dict(ck["discriminator"])
        self._gen_opt.load_state_dict(ck["gen_opt"])
        self._dis_opt.load_state_dict(ck["dis_opt"])
        self.counter = ck["counter"]

    def _save(self):
        os.makedirs(self.model_dir, exist_ok=True)
        torch.save({"generator": self.generator.state_dict(), "discriminator": self.discriminator.state_dict(),
                    "gen_opt": self._gen_opt.state_dict(), "dis_opt": self._dis_opt.state_dict(), "counter": self.counter},
                   self.checkpoint_path + ".pt")

    @staticmethod
    def _batches(dataset):
        if hasattr(dataset, "batches"):            # zoo TFDataset / FeatureSet
            for b in dataset.batches():
                yield b
        else:
            yield from dataset

    def train(self, input_fn, end_trigger):
        if isinstance(end_trigger, int):
            end_trigger = MaxIteration(end_trigger)
        if not isinstance(end_trigger, ZooTrigger):
            raise TypeError("end_trigger must be a ZooTrigger (e.g. MaxIteration(n))")
        period = self._discriminator_steps + self._generator_steps
        state = {"neval": self.counter + 1, "epoch": 1, "Loss": float("inf"), "score": 0.0}
        self.generator.train()
        self.discriminator.train()
        done = False
        while not done:
            seen = False
            for batch in self._batches(input_fn()):
                seen = True
                gen_in, real = batch[0], batch[1]
                to = lambda t: tuple(x.to(self.device) for x in t) if isinstance(t, (tuple, list)) else torch.as_tensor(t).to(self.device)
                gen_in, real = to(gen_in), to(real)
                d_phase = (self.counter % period) < self._discriminator_steps
                if d_phase:
                    with torch.no_grad():
                        fake = _call_maybe_with_counter(self.generator, self.counter, gen_in)
                    fake_d = _call_maybe_with_counter(self.discriminator, self.counter, fake, gen_in)
                    real_d = _call_maybe_with_counter(self.discriminator, self.counter, real, gen_in)
                    loss = _call_maybe_with_counter(self._discriminator_loss_fn, self.counter, real_d, fake_d)
                    self._dis_opt.zero_grad(set_to_none=True)
                    loss.backward()
                    self._dis_opt.step()
                else:
                    fake = _call_maybe_with_counter(self.generator, self.counter, gen_in)
                    fake_d = _call_maybe_with_counter(self.discriminator, self.counter, fake, gen_in)
                    loss = _call_maybe_with_counter(self._generator_loss_fn, self.counter, fake_d)
                    self._gen_opt.zero_grad(set_to_none=True)
                    loss.backward()                 # discriminator gradients are computed but never applied in this phase
                    self._gen_opt.step()
                self.counter += 1
                state["neval"], state["Loss"] = self.counter + 1, float(loss.detach())
                self.history.append((self.counter, "D" if d_phase else "G", state["Loss"]))
                if end_trigger(state):
                    done = True
                    break
            state["epoch"] += 1
            if not seen:
                raise ValueError("input_fn() produced no batches")
            if not done and end_trigger(state):
                done = True
        self._save()
        return self

    def generate(self, generator_inputs):
        self.generator.eval()
        with torch.no_grad():
            return self.generator(torch.as_tensor(generator_inputs).to(self.device)).cpu()
