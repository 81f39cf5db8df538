"""String -> optimizer / criterion / metric helpers (reference: pyzoo/zoo/pipeline/api/keras/utils.py `to_bigdl_*`)."""
from .metrics import to_metric
from .objectives import to_loss
from .optimizers import to_opt_config


def to_bigdl_optim_method(optimizer):
    return to_opt_config(optimizer)


def to_bigdl_criterion(criterion):
    return to_loss(criterion)


def to_bigdl_metric(metric, loss=None):
    return to_metric(metric, loss)


def to_bigdl_metrics(metrics, loss=None):
    return [to_metric(m, loss) for m in metrics]
