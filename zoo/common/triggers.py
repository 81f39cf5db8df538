"""Training triggers (end-of-training / checkpoint / validation conditions).

Reference: zoo/src/main/scala/.../common/ZooTrigger.scala:27-166 -- `EveryEpoch` (slice aware),
`SeveralIteration`, `MaxEpoch`, `MaxIteration`, `MaxScore`, `MinLoss`, `And`, `Or`.  A trigger is
called with the optimizer state table; the keys are the reference's: ``epoch`` (starts at 1),
``neval`` (next iteration number, starts at 1), ``Loss``, ``score``, plus ``numSlice`` /
``currentSlice`` for DISK_AND_DRAM feature sets.
"""
from __future__ import annotations


class ZooTrigger:
    def __call__(self, state: dict) -> bool:  # pragma: no cover - interface
        raise NotImplementedError

    def set_zoo_state(self, state):
        self.zoo_state = state
        return self

    def reset(self):
        pass


class EveryEpoch(ZooTrigger):
    """Fires once each time `epoch` advances (with numSlice > 1: once per full pass over all slices)."""

    def __init__(self):
        self.last_epoch = -1

    def __call__(self, state):
        ns = state.get("numSlice", 1)
        if ns > 1 and state.get("currentSlice", 0) % ns != 0:
            return False
        if self.last_epoch == -1:
            self.last_epoch = state["epoch"]
            return False
        if state["epoch"] == self.last_epoch:
            return False
        self.last_epoch = state["epoch"]
        return True

    def reset(self):
        self.last_epoch = -1


class SeveralIteration(ZooTrigger):
    def __init__(self, interval):
        if interval <= 0:
            raise ValueError("interval must be positive")
        self.interval = int(interval)

    def __call__(self, state):
        cur = state["neval"] - 1
        return cur != 0 and cur % self.interval == 0


class MaxEpoch(ZooTrigger):
    def __init__(self, max_epoch):
        self.max = int(max_epoch)

    def __call__(self, state):
        return state["epoch"] > self.max


class MaxIteration(ZooTrigger):
    def __init__(self, max_iteration):
        self.max = int(max_iteration)

    def __call__(self, state):
        return state["neval"] > self.max


class MaxScore(ZooTrigger):
    def __init__(self, max_score):
        self.max = float(max_score)

    def __call__(self, state):
        return state.get("score", float("-inf")) > self.max


class MinLoss(ZooTrigger):
    def __init__(self, min_loss):
        self.min = float(min_loss)

    def __call__(self, state):
        return state.get("Loss", float("inf")) < self.min


class And(ZooTrigger):
    def __init__(self, first, *others):
        self.triggers = [first, *others]

    def __call__(self, state):
        return all([t(state) for t in self.triggers])


class Or(ZooTrigger):
    def __init__(self, first, *others):
        self.triggers = [first, *others]

    def __call__(self, state):
        return any([t(state) for t in self.triggers])


# BigDL-style aliases used by the reference's Python examples (bigdl.optim.optimizer)
TriggerAnd, TriggerOr = And, Or
