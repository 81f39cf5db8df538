from .recommender import Recommender, UserItemFeature, UserItemPrediction  # noqa: F401
from .neuralcf import NeuralCF  # noqa: F401
from .wide_and_deep import WideAndDeep, ColumnFeatureInfo  # noqa: F401
from .session_recommender import SessionRecommender  # noqa: F401
from .utils import *  # noqa: F401,F403
