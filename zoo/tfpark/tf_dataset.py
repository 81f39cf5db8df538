"""TFDataset (reference: pyzoo/zoo/tfpark/tf_dataset.py:115-1179): the 11 constructors of the reference, producing a
batched, rank-shardable data set.  Here it wraps a `FeatureSet`; `batch_size` is the global training batch,
`batch_per_thread` the inference batch, exactly one of them (tf_dataset.py:160-178).  TFRecord files of tf.train.Example
records are read natively (`zoo.tfpark.tfrecord`); only `from_tf_data_dataset` needs TensorFlow itself."""
from __future__ import annotations

import numpy as np

from ..feature.common import FeatureSet


class TFDataset:
    def __init__(self, feature_set, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False, val_set=None):
        if batch_size > 0 and batch_per_thread > 0:
            raise ValueError("bath_size and batch_per_thread should not be set simultaneously")
        if batch_size <= 0 and batch_per_thread <= 0:
            batch_per_thread = 1
        self.feature_set, self.batch_size, self.batch_per_thread = feature_set, batch_size, batch_per_thread
        self.hard_code_batch_size, self.val_set = hard_code_batch_size, val_set
        self.has_batch = True

    # ---- constructors ------------------------------------------------------------------------------
    @staticmethod
    def from_ndarrays(tensors, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False, val_tensors=None, **kw):
        """tensors: x | (x, y) | nested tuples/lists of ndarrays."""
        def split(t):
            if isinstance(t, (tuple, list)) and len(t) == 2:
                return t[0], t[1]
            return t, None

        x, y = split(tensors)
        val = None
        if val_tensors is not None:
            vx, vy = split(val_tensors)
            val = FeatureSet.numpy(list(vx) if isinstance(vx, (tuple, list)) else vx, vy)
        fs = FeatureSet.numpy(list(x) if isinstance(x, (tuple, list)) else x, y)
        return TFDataset(fs, batch_size, batch_per_thread, hard_code_batch_size, val)

    @staticmethod
    def from_rdd(rdd, names=None, shapes=None, types=None, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False, val_rdd=None,
                 features=None, labels=None, **kw):
        """`rdd`: iterable of ndarray / (feature, label) records."""
        recs = list(rdd)
        if recs and isinstance(recs[0], (tuple, list)) and len(recs[0]) == 2:
            x, y = np.stack([np.asarray(r[0]) for r in recs]), np.stack([np.asarray(r[1]) for r in recs])
        else:
            x, y = np.stack([np.asarray(r) for r in recs]), None
        val = None
        if val_rdd is not None:
            val = TFDataset.from_rdd(val_rdd, batch_per_thread=1).feature_set
        return TFDataset(FeatureSet.numpy(x, y), batch_size, batch_per_thread, hard_code_batch_size, val)

    @staticmethod
    def from_feature_set(dataset, features=None, labels=None, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False,
                         validation_dataset=None, **kw):
        return TFDataset(dataset, batch_size, batch_per_thread, hard_code_batch_size, validation_dataset)

    @staticmethod
    def from_image_set(image_set, image, label=None, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False,
                       validation_image_set=None, **kw):
        val = validation_image_set.to_feature_set() if validation_image_set is not None else None
        return TFDataset(image_set.to_feature_set(), batch_size, batch_per_thread, hard_code_batch_size, val)

    @staticmethod
    def from_text_set(text_set, text, label=None, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False, validation_image_set=None, **kw):
        return TFDataset(text_set.to_feature_set(), batch_size, batch_per_thread, hard_code_batch_size)

    @staticmethod
    def from_dataframe(df, feature_cols, labels_cols=None, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False,
                       validation_df=None, **kw):
        def fs_of(d):
            x = np.stack([np.asarray(d[c].tolist(), dtype=np.float32) for c in feature_cols], 1) if len(feature_cols) > 1 \
                else np.asarray(d[feature_cols[0]].tolist(), dtype=np.float32)
            y = None if not labels_cols else np.asarray(d[labels_cols[0]].tolist())
            return FeatureSet.numpy(x.reshape(len(d), -1) if x.ndim == 1 else x, y)

        return TFDataset(fs_of(df), batch_size, batch_per_thread, hard_code_batch_size, fs_of(validation_df) if validation_df is not None else None)

    @staticmethod
    def _needs_tf(what):
        raise NotImplementedError(f"TFDataset.{what} reads TensorFlow data formats and needs the TensorFlow runtime, which is not part of "
                                  "zoo-b200; load the data into ndarrays and use from_ndarrays / from_feature_set")

    @staticmethod
    def from_tfrecord_file(sc=None, file_path=None, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False, validation_file_path=None,
                           features=None, labels=None, image_decoder=None, **kw):
        """TFRecord files of tf.train.Example records (read natively, `zoo.tfpark.tfrecord`).  The reference returns the serialized
        strings for a TF parsing graph; here `features` / `labels` name the Example features that become the dataset's tensors
        (all features except `labels` when `features` is None)."""
        from .tfrecord import examples_to_arrays, is_numeric_feature, read_examples

        if file_path is None and isinstance(sc, str):      # from_tfrecord_file(path, ...) without a SparkContext
            sc, file_path = None, sc

        def load(path):
            exs = list(read_examples(path))
            if not exs:
                raise ValueError(f"no records in {path}")
            lab = [] if labels is None else ([labels] if isinstance(labels, str) else list(labels))
            # default feature set: everything that can be a tensor (numbers, encoded images) except the labels
            usable = lambda k: image_decoder is not None or is_numeric_feature(exs[0][k])
            feats = features if features is not None else sorted(k for k in exs[0] if k not in lab and usable(k))
            return examples_to_arrays(exs, feats, labels, image_decoder)

        x, y = load(file_path)
        val = None
        if validation_file_path is not None:
            vx, vy = load(validation_file_path)
            val = FeatureSet.numpy(vx, vy)
        return TFDataset(FeatureSet.numpy(x, y), batch_size, batch_per_thread, hard_code_batch_size, val)

    @staticmethod
    def from_tf_data_dataset(dataset, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False, validation_dataset=None, **kw):
        """A `tf.data.Dataset` in the reference (pyzoo/zoo/tfpark/tf_dataset.py:560-640: the dataset graph is shipped to the executors and
        iterated there).  There is no TensorFlow runtime here, so what is accepted is the thing a tf.data pipeline IS to its consumer:
        any iterable of elements, an element being an array, or a (features, labels) pair whose sides are arrays or tuples of arrays.
        A real `tf.data.Dataset` object cannot exist in this environment; one that does (a TF install next to zoo) is iterated through
        its `as_numpy_iterator()`."""
        def collect(ds):
            it = ds.as_numpy_iterator() if hasattr(ds, "as_numpy_iterator") else iter(ds)
            elems = list(it)
            if not elems:
                raise ValueError("from_tf_data_dataset: the dataset is empty")
            pair = isinstance(elems[0], (tuple, list)) and len(elems[0]) == 2
            side = lambda vals: ([np.stack([np.asarray(v[k]) for v in vals]) for k in range(len(vals[0]))] if isinstance(vals[0], (tuple, list))
                                 else np.stack([np.asarray(v) for v in vals]))
            if pair:
                return side([e[0] for e in elems]), side([e[1] for e in elems])
            return side(elems), None

        x, y = collect(dataset)
        val = None
        if validation_dataset is not None:
            vx, vy = collect(validation_dataset)
            val = FeatureSet.numpy(vx, vy)
        return TFDataset(FeatureSet.numpy(x, y), batch_size, batch_per_thread, hard_code_batch_size, val)

    @staticmethod
    def from_string_rdd(string_rdd, batch_size=-1, batch_per_thread=-1, hard_code_batch_size=False, validation_string_rdd=None, **kw):
        return TFDataset(FeatureSet.rdd([np.frombuffer(s.encode() if isinstance(s, str) else s, np.uint8) for s in string_rdd]),
                         batch_size, batch_per_thread, hard_code_batch_size)

    from_bytes_rdd = from_string_rdd

    # ---- accessors ------------------------------------------------------------------------------------
    def get_training_data(self):
        return self.feature_set

    def get_validation_data(self):
        return self.val_set

    def get_prediction_data(self):
        return self.feature_set

    def get_evaluation_data(self):
        return self.feature_set

    def get_num_partitions(self):
        from ..common.nncontext import get_context

        return get_context().world_size

    def __len__(self):
        return self.feature_set.size()


# concrete dataset class names of the reference (tf_dataset.py:600-1179); all are TFDataset here
TFNdarrayDataset = TFFeatureDataset = TFBytesDataset = TFTextDataset = TFImageDataset = DataFrameDataset = TFDataDataset = TFDataset


class TensorMeta:
    def __init__(self, dtype, name=None, shape=None):
        self.dtype, self.name, self.shape = dtype, name, shape


class MergeFeatureLabelImagePreprocessing:
    def __call__(self, f):
        return f


MergeFeatureLabelFeatureTransformer = MergeFeatureLabelImagePreprocessing
