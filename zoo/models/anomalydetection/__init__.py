from .anomaly_detector import AnomalyDetector, FeatureLableIndex  # noqa: F401
