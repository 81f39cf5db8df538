from .forecast import Forecaster, LSTMForecaster, MTNetForecaster, TCNForecaster  # noqa: F401
