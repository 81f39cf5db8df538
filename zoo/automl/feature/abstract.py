"""BaseFeatureTransformer contract (reference: pyzoo/zoo/automl/feature/abstract.py)."""
from abc import ABC, abstractmethod


class BaseFeatureTransformer(ABC):
    check_optional_config = False

    @abstractmethod
    def fit_transform(self, input_df, **config):
        pass

    @abstractmethod
    def transform(self, input_df, is_train=True):
        pass

    @abstractmethod
    def save(self, file_path):
        pass

    @abstractmethod
    def restore(self, **config):
        pass
