from .forecast import AutoTSTrainer, TSPipeline  # noqa: F401
