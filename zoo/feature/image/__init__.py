from ..common import ChainedPreprocessing  # noqa: F401
from .imagePreprocessing import *  # noqa: F401,F403
from .imageset import DistributedImageSet, ImageFeature, ImageSet, LocalImageSet  # noqa: F401
