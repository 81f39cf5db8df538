"""zoo.pipeline.api.net: TorchNet / TorchCriterion / Net loaders / GraphNet / TFNet (reference: pyzoo/zoo/pipeline/api/net/*)."""
from .graph_net import GraphNet  # noqa: F401
from .net_load import Net  # noqa: F401
from .tfnet import GraphRunner, TFNet, TFNetForInference  # noqa: F401
from .torch_criterion import LossWrapperModule, TorchCriterion  # noqa: F401
from .torch_net import TorchNet  # noqa: F401
