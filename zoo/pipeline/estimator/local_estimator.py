"""LocalEstimator (reference: zoo/.../pipeline/estimator/LocalEstimator.scala:39-309): single-process training on
plain in-memory samples -- the reference clones the model `threadNum` times, splits each batch over the clones, sums
their gradients slice-parallel and takes one optimizer step.  One GPU replaces the thread replicas, so this
is the engine with world_size 1 behind the reference's `fit(trainData, testData, transformer, batchSize, epoch)`."""
from __future__ import annotations

import numpy as np

from ...common.triggers import MaxEpoch
from ...feature.common import FeatureSet, Sample
from ..engine import TrainingEngine


class LocalEstimator:
    def __init__(self, model, criterion, optmize_method, validations=None, thread_num=1):
        self.model, self.criterion, self.validations, self.thread_num = model, criterion, list(validations or []), thread_num
        self.engine = TrainingEngine(model, optmize_method)

    @staticmethod
    def _to_set(data, transformer=None):
        recs = [transformer(d) if transformer is not None else d for d in data]
        if recs and isinstance(recs[0], Sample):
            return FeatureSet.sample_rdd(recs)
        x = np.stack([np.asarray(r[0]) for r in recs])
        y = np.stack([np.asarray(r[1]) for r in recs])
        return FeatureSet.numpy(x, y)

    def fit(self, train_data, test_data=None, transformer=None, batch_size=32, epoch=1):
        train = self._to_set(train_data, transformer)
        test = self._to_set(test_data, transformer) if test_data is not None else None
        end = MaxEpoch(self.engine.state["epoch"] + epoch - 1)
        from ...common.triggers import EveryEpoch

        self.engine.train(train, self.criterion, end, EveryEpoch() if test is not None and self.validations else None, test,
                          self.validations or None, batch_size)
        return self

    def validate(self, data, transformer=None, batch_size=32):
        return self.engine.evaluate(self._to_set(data, transformer), self.validations or ["loss"], batch_size, self.criterion)
