from .nncontext import *  # noqa: F401,F403
from .triggers import (  # noqa: F401
    ZooTrigger, EveryEpoch, SeveralIteration, MaxEpoch, MaxIteration, MaxScore, MinLoss, And, Or, TriggerAnd, TriggerOr,
)
