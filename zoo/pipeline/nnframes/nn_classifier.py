"""NNFrames: DataFrame-in / DataFrame-out estimators and transformers.

Reference: pyzoo/zoo/pipeline/nnframes/nn_classifier.py:135-600 and
zoo/src/main/scala/.../pipeline/nnframes/{NNEstimator,NNClassifier}.scala -- Spark-ML `Estimator` /
`Transformer` stages over Spark DataFrames: `NNEstimator(model, criterion, feature_preprocessing,
label_preprocessing).setBatchSize(..).setMaxEpoch(..)....fit(df)` returns an `NNModel` whose `transform(df)`
appends a prediction column; `NNClassifier` / `NNClassifierModel` add arg-max (1-based) / 0.5-threshold
prediction; models persist as a directory (`metadata/` JSON with `tensorDataType` / `modelType`, `module` +
`weight` | `torchscript.pt`, `samplePreprocessing`).

There is no Spark here: a DataFrame is a pandas DataFrame (or a pyarrow Table, converted).  Setter names,
defaults (batch size 1, max epoch 50, learning rate 1e-3, SGD), validation / checkpoint / summary / clipping
plumbing, prediction semantics and the on-disk layout are the reference's; training runs on the engine
(`zoo.pipeline.engine`), so a multi-GPU launch shards the rows over the ranks like `FeatureSet.rdd` coalesces
to one partition per node.
"""
from __future__ import annotations

import json
import os
import pickle
import time

import numpy as np
import torch
import torch.nn as nn

from ...common.triggers import EveryEpoch, MaxEpoch
from ...feature.common import (ChainedPreprocessing, FeatureLabelPreprocessing, FeatureSet, Preprocessing, ScalarToTensor,
                               SeqToTensor, TensorToSample)
from ..api.keras.optimizers import SGD
from ..engine import TrainingEngine


def _to_pandas(df):
    import pandas as pd

    if isinstance(df, pd.DataFrame):
        return df
    if hasattr(df, "to_pandas"):
        return df.to_pandas()
    raise TypeError(f"expected a pandas DataFrame (or pyarrow Table), got {type(df)}")


def _default_preprocessing(p):
    if p is None:
        return SeqToTensor()
    if isinstance(p, Preprocessing):
        return p
    if isinstance(p, (list, tuple)):        # a size list -> SeqToTensor(size)
        return SeqToTensor(list(p))
    raise TypeError(f"cannot interpret {p!r} as a Preprocessing")


def _cell_to_array(v, prep):
    out = prep(v)
    if hasattr(out, "features"):     # a Sample
        out = out.features[0]
    return np.asarray(out)


class _Params:
    """Spark-ML style param accessors shared by the estimator and the model."""

    def _init_params(self):
        self.featuresCol, self.labelCol, self.predictionCol, self.batchSize = "features", "label", "prediction", 1

    def setFeaturesCol(self, v):
        self.featuresCol = v
        return self

    def getFeaturesCol(self):
        return self.featuresCol

    def setLabelCol(self, v):
        self.labelCol = v
        return self

    def getLabelCol(self):
        return self.labelCol

    def setPredictionCol(self, v):
        self.predictionCol = v
        return self

    def getPredictionCol(self):
        return self.predictionCol

    def setBatchSize(self, v):
        self.batchSize = int(v)
        return self

    def getBatchSize(self):
        return self.batchSize

    def setSamplePreprocessing(self, v):
        self.samplePreprocessing = v
        return self

    def getSamplePreprocessing(self):
        return self.samplePreprocessing


class NNModel(_Params):
    """Transformer: appends `predictionCol` (the model output per row) to a DataFrame."""
    model_type_name = "NNModel"

    def __init__(self, model, feature_preprocessing=None, jvalue=None, bigdl_type="float"):
        self._init_params()
        self.model = model
        self.feature_preprocessing = _default_preprocessing(feature_preprocessing)
        self.samplePreprocessing = ChainedPreprocessing([self.feature_preprocessing, TensorToSample()])

    def _features(self, df):
        col = df[self.featuresCol]
        return np.stack([_cell_to_array(v, self.feature_preprocessing) for v in col]).astype(np.float32)

    def _predict_array(self, df):
        x = self._features(df)
        eng = TrainingEngine(self.model)
        return eng.predict(FeatureSet.numpy(x), batch_size=max(self.batchSize, 1) * 64)

    def _output_to_prediction(self, out):
        return [row.tolist() for row in out.reshape(len(out), -1)]

    def transform(self, df):
        df = _to_pandas(df)
        out = df.copy()
        out[self.predictionCol] = self._output_to_prediction(np.asarray(self._predict_array(df)))
        return out

    # ---- ML persistence (NNEstimator.scala:805-918) ---------------------------------------------------
    def save(self, path, over_write=True):
        if os.path.exists(path) and not over_write:
            raise IOError(f"{path} exists")
        os.makedirs(os.path.join(path, "metadata"), exist_ok=True)
        from ..api.net.torch_net import TorchNet

        if isinstance(self.model, TorchNet):
            model_type = "TorchNet"
            try:
                self.model.savePytorch(os.path.join(path, "torchscript.pt"))
            except Exception:        # not scriptable: fall back to a pickled module next to it
                torch.save(self.model, os.path.join(path, "module"))
                model_type = "default"
        else:
            model_type = "default"
            torch.save(self.model, os.path.join(path, "module"))
            torch.save(self.model.state_dict(), os.path.join(path, "weight"))
        with open(os.path.join(path, "samplePreprocessing"), "wb") as f:
            pickle.dump(self.feature_preprocessing, f)
        meta = {"class": f"zoo.pipeline.nnframes.{type(self).__name__}", "timestamp": int(time.time() * 1000), "uid": f"{type(self).__name__}_{id(self):x}",
                "paramMap": {"featuresCol": self.featuresCol, "predictionCol": self.predictionCol, "batchSize": self.batchSize,
                             **({"threshold": self.threshold} if hasattr(self, "threshold") else {})},
                "tensorDataType": "TensorFloat", "modelType": model_type}
        with open(os.path.join(path, "metadata", "part-00000"), "w") as f:
            f.write(json.dumps(meta) + "\n")
        return self

    def write(self):
        return _Writer(self)

    @classmethod
    def load(cls, path):
        """A directory written by `save` here, or by the reference's NNModel writer (NNEstimator.scala:805-918: `metadata/part-00000`
        JSON, `module` + `weight` in BigDL's format, `samplePreprocessing` Java-serialized -- the latter cannot be decoded outside
        a JVM, so such a model gets the default array-to-tensor preprocessing)."""
        with open(os.path.join(path, "metadata", "part-00000")) as f:
            meta = json.loads(f.readline())
        module_file = os.path.join(path, "module")
        with open(module_file, "rb") as f:
            head = f.read(2) if os.path.exists(module_file) else b""
        prep = None
        if meta.get("modelType") == "TorchNet":
            from ..api.net.torch_net import TorchNet

            model = TorchNet(os.path.join(path, "torchscript.pt"))
        elif head in (b"PK", b"\x80\x02", b"\x80\x03", b"\x80\x04", b"\x80\x05"):
            model = torch.load(module_file, map_location="cpu", weights_only=False)
        else:                                                      # the reference's own persistence
            from ...models.bigdl import load_bigdl

            weight_file = os.path.join(path, "weight")
            model = load_bigdl(module_file, weight_file if os.path.exists(weight_file) else None)
        try:
            with open(os.path.join(path, "samplePreprocessing"), "rb") as f:
                if f.read(2) != b"\xac\xed":                      # 0xACED: a Java object stream
                    f.seek(0)
                    prep = pickle.load(f)
        except (OSError, pickle.UnpicklingError, EOFError):
            prep = None
        klass = NNClassifierModel if meta["class"].endswith("NNClassifierModel") else NNModel
        m = klass(model, prep)
        pm = meta["paramMap"]
        m.setFeaturesCol(pm["featuresCol"]).setPredictionCol(pm["predictionCol"]).setBatchSize(pm["batchSize"])
        if "threshold" in pm and hasattr(m, "setThreshold"):
            m.setThreshold(pm["threshold"])
        return m


class _Writer:
    def __init__(self, m):
        self.m, self._ow = m, False

    def overwrite(self):
        self._ow = True
        return self

    def save(self, path):
        return self.m.save(path, over_write=self._ow)


class NNClassifierModel(NNModel):
    """Prediction = arg-max class as a 1-based double, or 1.0 / 0.0 by threshold when the model has one output
    (NNClassifier.scala:167-210)."""

    def __init__(self, model, feature_preprocessing=None, jvalue=None, bigdl_type="float"):
        super().__init__(model, feature_preprocessing)
        self.threshold = 0.5

    def setThreshold(self, v):
        self.threshold = float(v)
        return self

    def getThreshold(self):
        return self.threshold

    def _output_to_prediction(self, out):
        out = out.reshape(len(out), -1)
        if out.shape[1] == 1:
            return (out[:, 0] > self.threshold).astype(np.float64)
        return (out.argmax(1) + 1).astype(np.float64)


class NNEstimator(_Params):
    model_class = NNModel

    def __init__(self, model, criterion, feature_preprocessing=None, label_preprocessing=None, jvalue=None, bigdl_type="float"):
        self._init_params()
        self.model, self.criterion = model, criterion
        self.feature_preprocessing = _default_preprocessing(feature_preprocessing)
        self.label_preprocessing = _default_preprocessing(label_preprocessing)
        self.samplePreprocessing = FeatureLabelPreprocessing(self.feature_preprocessing, self.label_preprocessing)
        self.maxEpoch, self.learningRate, self.learningRateDecay = 50, 1e-3, 0.0
        self.optimMethod = None
        self.endWhen = None
        self.cachingSample, self.dataCacheLevel = True, ("DRAM", None)
        self.train_summary = self.validation_summary = None
        self.validation = None
        self.clip = None
        self.checkpoint = None
        self._engine = None

    # ---- setters (nn_classifier.py:224-437) --------------------------------------------------------------
    def setSamplePreprocessing(self, val):
        self.samplePreprocessing = val
        return self

    def setMaxEpoch(self, val):
        self.maxEpoch = int(val)
        return self

    def getMaxEpoch(self):
        return self.maxEpoch

    def setEndWhen(self, trigger):
        self.endWhen = trigger
        return self

    def getEndWhen(self):
        return self.endWhen

    def setDataCacheLevel(self, level, numSlice=None):
        if level == "DISK_AND_DRAM" and not numSlice:
            raise ValueError("numSlice is required for DISK_AND_DRAM")
        self.dataCacheLevel = (level, numSlice)
        return self

    def getDataCacheLevel(self):
        return self.dataCacheLevel

    def setLearningRate(self, val):
        self.learningRate = float(val)
        return self

    def getLearningRate(self):
        return self.learningRate

    def setLearningRateDecay(self, val):
        self.learningRateDecay = float(val)
        return self

    def getLearningRateDecay(self):
        return self.learningRateDecay

    def setOptimMethod(self, val):
        self.optimMethod = val
        return self

    def getOptimMethod(self):
        return self.optimMethod

    def setCachingSample(self, val):
        self.cachingSample = bool(val)
        return self

    def isCachingSample(self):
        return self.cachingSample

    def setTrainSummary(self, val):
        self.train_summary = val
        return self

    def getTrainSummary(self):
        return self.train_summary

    def setValidationSummary(self, val):
        self.validation_summary = val
        return self

    def getValidationSummary(self):
        return self.validation_summary

    def setValidation(self, trigger, val_df, val_method, batch_size):
        self.validation = (trigger, val_df, list(val_method), int(batch_size))
        return self

    def getValidation(self):
        return self.validation

    def clearGradientClipping(self):
        self.clip = None
        return self

    def setConstantGradientClipping(self, min, max):
        self.clip = ("const", float(min), float(max))
        return self

    def setGradientClippingByL2Norm(self, clip_norm):
        self.clip = ("l2", float(clip_norm))
        return self

    def setCheckpoint(self, path, trigger, isOverWrite=True):
        self.checkpoint = (path, trigger, isOverWrite)
        return self

    def getCheckpoint(self):
        return self.checkpoint

    # ---- fit ---------------------------------------------------------------------------------------------
    def _label_array(self, col):
        return np.stack([_cell_to_array(v, self.label_preprocessing) for v in col])

    def _to_set(self, df):
        df = _to_pandas(df)
        x = np.stack([_cell_to_array(v, self.feature_preprocessing) for v in df[self.featuresCol]]).astype(np.float32)
        y = self._label_array(df[self.labelCol])
        level, n = self.dataCacheLevel
        mem = f"DISK_AND_DRAM({n})" if level == "DISK_AND_DRAM" else level
        return FeatureSet.numpy(x, y, memory_type=mem)

    def fit(self, df):
        opt = self.optimMethod if self.optimMethod is not None else SGD(learningrate=self.learningRate, learningrate_decay=self.learningRateDecay)
        if self._engine is None or self._engine.optim_methods is not opt:
            self._engine = TrainingEngine(self.model, opt)
        eng = self._engine
        if self.clip is not None:
            (eng.set_constant_gradient_clipping if self.clip[0] == "const" else eng.set_gradient_clipping_by_l2_norm)(*self.clip[1:])
        if self.train_summary is not None:
            eng.train_summary = self.train_summary
        if self.validation_summary is not None:
            eng.val_summary = self.validation_summary
        trig, val_set, val_methods = None, None, None
        if self.validation is not None:
            trig, vdf, val_methods, _ = self.validation
            val_set = self._to_set(vdf)
        if self.checkpoint is not None:
            eng.set_checkpoint(self.checkpoint[0], self.checkpoint[2])
            trig = trig or self.checkpoint[1]
        end = self.endWhen or MaxEpoch(eng.state["epoch"] + self.maxEpoch - 1)
        eng.train(self._to_set(df), self.criterion, end, trig or (EveryEpoch() if val_set is not None else None), val_set, val_methods,
                  batch_size=self.batchSize)
        return self._create_model()

    def _create_model(self):
        m = self.model_class(self.model, self.feature_preprocessing)
        m.setFeaturesCol(self.featuresCol).setPredictionCol(self.predictionCol).setBatchSize(self.batchSize)
        return m


class NNClassifier(NNEstimator):
    """Labels are class indices; with a zoo classification criterion they are 1-based, as in the reference."""
    model_class = NNClassifierModel

    def __init__(self, model, criterion, feature_preprocessing=None, jvalue=None, bigdl_type="float"):
        super().__init__(model, criterion, feature_preprocessing, ScalarToTensor())

    def _label_array(self, col):
        return np.asarray([float(np.asarray(v).reshape(-1)[0]) for v in col], dtype=np.float32)


# Spark-ML param mixin names of the reference (nn_classifier.py:29-132); here one mixin carries them all
HasBatchSize = HasSamplePreprocessing = HasOptimMethod = _Params


class HasThreshold:
    def setThreshold(self, v):
        self.threshold = float(v)
        return self

    def getThreshold(self):
        return self.threshold
