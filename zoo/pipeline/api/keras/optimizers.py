"""Optimizers of the keras-style API, expressed as configurations of the fused kernels.

Reference: zoo/src/main/scala/.../pipeline/api/keras/optimizers/Adam.scala:38-135 (Adam with a
learning-rate schedule, `PolyEpochDecay`), AdamWeightDecay.scala:40-156 (BERT Adam with
linear/cosine/constant warm-up), and the BigDL OptimMethods the reference wires through
`compile("adam" | "sgd" | "rmsprop" | "adagrad" | "adadelta" | "adamax")` and per-sub-module
Estimator maps (Ftrl/Adagrad for WideAndDeep, CensusWideAndDeep.scala:122-132).
"""
from __future__ import annotations

import math

from ....parallel.flat import OptConfig


class Schedule:
    def __call__(self, step, epoch, base_lr):
        return base_lr


class Default(Schedule):
    """BigDL SGD.Default: lr / (1 + step * decay)."""

    def __init__(self, decay=0.0):
        self.decay = decay

    def __call__(self, step, epoch, base_lr):
        return base_lr / (1.0 + max(step - 1, 0) * self.decay)


class Fixed(Schedule):
    """zoo `Optim.Fixed` (common/Optim.scala)."""


class PolyEpochDecay(Schedule):
    """lr * (1 - epoch/max_epochs)^power (Adam.scala:135)."""

    def __init__(self, power, max_epochs):
        self.power, self.max_epochs = power, max_epochs

    def __call__(self, step, epoch, base_lr):
        return base_lr * (1.0 - min(epoch, self.max_epochs) / float(self.max_epochs)) ** self.power


class Poly(Schedule):
    def __init__(self, power, max_iteration):
        self.power, self.max_iteration = power, max_iteration

    def __call__(self, step, epoch, base_lr):
        return base_lr * (1.0 - min(step, self.max_iteration) / float(self.max_iteration)) ** self.power


class Step(Schedule):
    def __init__(self, step_size, gamma):
        self.step_size, self.gamma = step_size, gamma

    def __call__(self, step, epoch, base_lr):
        return base_lr * self.gamma ** (step // self.step_size)


class EpochStep(Schedule):
    def __init__(self, step_size, gamma):
        self.step_size, self.gamma = step_size, gamma

    def __call__(self, step, epoch, base_lr):
        return base_lr * self.gamma ** (epoch // self.step_size)


class Exponential(Schedule):
    def __init__(self, decay_step, decay_rate, stair_case=False):
        self.decay_step, self.decay_rate, self.stair_case = decay_step, decay_rate, stair_case

    def __call__(self, step, epoch, base_lr):
        e = step / float(self.decay_step)
        return base_lr * self.decay_rate ** (math.floor(e) if self.stair_case else e)


class Warmup(Schedule):
    """Linear warm-up by `delta` per iteration (BigDL SGD.Warmup), e.g. ResNet-50's 0.1 -> 3.2."""

    def __init__(self, delta, warmup_iteration=None):
        self.delta, self.warmup_iteration = delta, warmup_iteration

    def __call__(self, step, epoch, base_lr):
        s = step if self.warmup_iteration is None else min(step, self.warmup_iteration)
        return base_lr + self.delta * s


class SequentialSchedule(Schedule):
    def __init__(self, iteration_per_epoch=1):
        self.parts = []

    def add(self, schedule, max_iteration):
        self.parts.append((schedule, max_iteration))
        return self

    def __call__(self, step, epoch, base_lr):
        off = 0
        for sch, n in self.parts:
            if step < off + n or (sch, n) == self.parts[-1]:
                return sch(step - off, epoch, base_lr)
            off += n
        return base_lr


def _warmup_linear(x, warmup):
    return x / warmup if x < warmup else 1.0 - x


def _warmup_cosine(x, warmup):
    return x / warmup if x < warmup else 0.5 * (1.0 + math.cos(math.pi * x))


def _warmup_constant(x, warmup):
    return x / warmup if x < warmup else 1.0


class OptimMethod:
    """Common face: `.to_config()` -> OptConfig for the fused kernels."""

    def to_config(self) -> OptConfig:  # pragma: no cover
        raise NotImplementedError


class Adam(OptimMethod):
    def __init__(self, lr=1e-3, beta_1=0.9, beta_2=0.999, epsilon=1e-8, decay=0.0, schedule=None):
        self.lr, self.beta_1, self.beta_2, self.epsilon, self.decay = lr, beta_1, beta_2, epsilon, decay
        self.schedule = schedule if schedule is not None else Default(decay)

    def to_config(self):
        return OptConfig("adam_zoo", self.lr, self.beta_1, self.beta_2, self.epsilon, schedule=self.schedule)


class AdamWeightDecay(OptimMethod):
    def __init__(self, lr=1e-3, warmup_portion=-1.0, total=-1, schedule="linear", beta1=0.9, beta2=0.999,
                 epsilon=1e-6, weight_decay=0.01):
        self.lr, self.warmup_portion, self.total = lr, warmup_portion, total
        self.beta1, self.beta2, self.epsilon, self.weight_decay = beta1, beta2, epsilon, weight_decay
        self.warm = {"linear": _warmup_linear, "cosine": _warmup_cosine, "constant": _warmup_constant}[schedule]

    def to_config(self):
        def sch(step, epoch, base_lr):
            if self.total == -1:
                return base_lr
            x = max(step - 1, 0) / float(self.total)
            f = self.warm(x, self.warmup_portion)
            return base_lr * f * f  # the reference applies the warm-up factor twice (AdamWeightDecay.scala:113-117)

        return OptConfig("adamw_bert", self.lr, self.beta1, self.beta2, self.epsilon, self.weight_decay, schedule=sch)


class SGD(OptimMethod):
    def __init__(self, learningrate=1e-3, learningrate_decay=0.0, weightdecay=0.0, momentum=0.0, dampening=None,
                 nesterov=False, leaningrate_schedule=None, learningrate_schedule=None):
        self.lr, self.decay, self.wd, self.momentum = learningrate, learningrate_decay, weightdecay, momentum
        self.dampening = momentum if dampening is None else dampening  # BigDL default: dampening = momentum... (0 w/ nesterov)
        if dampening is None:
            self.dampening = 0.0
        self.nesterov = nesterov
        self.schedule = learningrate_schedule or leaningrate_schedule or Default(learningrate_decay)

    def to_config(self):
        return OptConfig("sgd", self.lr, self.momentum, 0.0, 0.0, self.wd, x1=self.dampening,
                         x2=1.0 if self.nesterov else 0.0, schedule=self.schedule)


class Adagrad(OptimMethod):
    def __init__(self, learningrate=1e-3, learningrate_decay=0.0, weightdecay=0.0):
        self.lr, self.decay, self.wd = learningrate, learningrate_decay, weightdecay

    def to_config(self):
        return OptConfig("adagrad", self.lr, 0.0, 0.0, 1e-10, self.wd, schedule=Default(self.decay))


class RMSprop(OptimMethod):
    def __init__(self, learningrate=1e-2, learningrate_decay=0.0, decayrate=0.99, epsilon=1e-8):
        self.lr, self.decay, self.rho, self.eps = learningrate, learningrate_decay, decayrate, epsilon

    def to_config(self):
        return OptConfig("rmsprop", self.lr, 0.0, self.rho, self.eps, schedule=Default(self.decay))


class Ftrl(OptimMethod):
    def __init__(self, learningrate=1e-3, learningrate_power=-0.5, initial_accumulator_value=0.1,
                 l1_regularization_strength=0.0, l2_regularization_strength=0.0, l2_shrinkage_regularization_strength=0.0):
        self.lr, self.lrp, self.init = learningrate, learningrate_power, initial_accumulator_value
        self.l1, self.l2 = l1_regularization_strength, l2_regularization_strength

    def to_config(self):
        return OptConfig("ftrl", self.lr, x1=self.l1, x2=self.l2, x3=self.lrp, init_accum=self.init)


def to_opt_config(opt) -> OptConfig:
    """str | OptimMethod | OptConfig | torch.optim.Optimizer -> OptConfig (KerasUtils.toBigDLOptimMethod)."""
    import torch

    if isinstance(opt, OptConfig):
        return opt
    if isinstance(opt, OptimMethod):
        return opt.to_config()
    if isinstance(opt, torch.optim.Optimizer):
        return OptConfig.from_torch(opt)
    if isinstance(opt, type) and issubclass(opt, torch.optim.Optimizer):
        # an optimizer CLASS: its torch defaults (lr 1e-3 where torch requires one)
        probe = torch.nn.Parameter(torch.zeros(1))
        try:
            return OptConfig.from_torch(opt([probe]))
        except TypeError:
            return OptConfig.from_torch(opt([probe], lr=1e-3))
    if isinstance(opt, type) and issubclass(opt, OptimMethod):
        return opt().to_config()
    if isinstance(opt, str):
        k = opt.lower()
        table = {"sgd": lambda: SGD(learningrate=0.01), "adam": Adam, "rmsprop": RMSprop, "adagrad": lambda: Adagrad(0.01),
                 "ftrl": Ftrl, "adamweightdecay": AdamWeightDecay}
        if k in table:
            return table[k]().to_config()
        raise ValueError(f"unsupported optimizer {opt!r}")
    raise TypeError(f"cannot interpret {opt!r} as an optimizer")
