"""Pipeline contract (reference: pyzoo/zoo/automl/pipeline/abstract.py)."""
from abc import ABC, abstractmethod


class Pipeline(ABC):
    @abstractmethod
    def fit(self, input_df, validation_df=None, **kw):
        pass

    @abstractmethod
    def evaluate(self, input_df, metrics=("mse",)):
        pass

    @abstractmethod
    def predict(self, input_df):
        pass
