"""BaseModel contract (reference: pyzoo/zoo/automl/model/abstract.py:20-105)."""
from __future__ import annotations

from abc import ABC, abstractmethod


class BaseModel(ABC):
    check_optional_config = False
    future_seq_len = None

    @abstractmethod
    def fit_eval(self, x, y, validation_data=None, mc=False, verbose=0, **config):
        """One call = `epochs` more epochs; returns the reward metric on the validation (or training) data."""

    @abstractmethod
    def evaluate(self, x, y, metric=None):
        pass

    @abstractmethod
    def predict(self, x, mc=False):
        pass

    @abstractmethod
    def save(self, model_path, config_path):
        pass

    @abstractmethod
    def restore(self, model_path, **config):
        pass

    def _get_required_parameters(self):
        return set()

    def _get_optional_parameters(self):
        return set()

    def _check_config(self, **config):
        missing = self._get_required_parameters() - set(config)
        if missing:
            raise ValueError(f"Missing required parameters in configuration: {sorted(missing)}")
        if self.check_optional_config:
            unset = self._get_optional_parameters() - set(config)
            if unset:
                raise ValueError(f"Missing optional parameters in configuration: {sorted(unset)}")
        return True
