from .zoo_model import ZooModel, KerasZooModel  # noqa: F401
from .ranker import Ranker  # noqa: F401
