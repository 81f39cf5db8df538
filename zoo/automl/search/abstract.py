"""SearchEngine / TrialOutput contracts (reference: pyzoo/zoo/automl/search/abstract.py:20-66)."""
from __future__ import annotations

from abc import ABC, abstractmethod


class TrialOutput:
    def __init__(self, config, model_path):
        self.config, self.model_path = config, model_path


class SearchEngine(ABC):
    @abstractmethod
    def run(self):
        pass

    @abstractmethod
    def get_best_trials(self, k):
        pass


from ..config.recipe import GridSearch, RandomSample  # noqa: E402,F401


class GoodError(Exception):
    pass


class BayersianOpt:
    """Marker for a Bayesian-optimisation search (see BayesRecipe / RayTuneSearchEngine)."""
