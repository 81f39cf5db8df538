"""ImageSet / ImageFeature (reference: pyzoo/zoo/feature/image/imageset.py:21-211, zoo/.../feature/image/ImageSet.scala).

An `ImageFeature` is a dict with the reference's keys: `uri`, `bytes` (encoded file), `mat` (HWC BGR ndarray, the
OpenCV Mat), `floats` / `imageTensor` (CHW float tensor), `label`, `originalSize`, `predict`, `sample`.
`ImageSet.read(path)` reads a file / directory (optionally one sub-directory per class: `with_label=True`),
`transform` applies a preprocessing chain lazily per feature, `get_image / get_label / get_predict` collect results.
There is no Spark: Local and Distributed sets are the same in-memory list; a multi-GPU job shards it by rank when it is
turned into a FeatureSet."""
from __future__ import annotations

import glob
import os

import numpy as np

IMG_EXT = (".jpg", ".jpeg", ".png", ".bmp", ".ppm", ".pgm", ".tif", ".tiff", ".webp")


class ImageFeature(dict):
    def get_tensor(self, key=None):
        for k in ([key] if key else ["imageTensor", "floats"]):
            if k in self:
                return np.asarray(self[k], dtype=np.float32)
        m = np.asarray(self["mat"], dtype=np.float32)
        return m.transpose(2, 0, 1) if m.ndim == 3 else m[None]

    def get_label(self):
        return self.get("label")

    def get_image(self, key="floats", to_chw=True):
        if key in self:
            return np.asarray(self[key])
        m = np.asarray(self["mat"], dtype=np.float32)
        return m.transpose(2, 0, 1) if to_chw else m


class ImageSet:
    def __init__(self, features=None, label_map=None, transformer=None):
        self.features = list(features or [])
        self._label_map = label_map
        self.transformer = transformer

    # ---- constructors ----------------------------------------------------------------------
    @classmethod
    def read(cls, path, sc=None, min_partitions=1, resize_height=-1, resize_width=-1, image_codec=-1, with_label=False, one_based_label=True,
             bigdl_type="float"):
        import cv2

        path = path[len("file://"):] if str(path).startswith("file://") else str(path)
        files, labels, label_map = [], [], None
        if with_label:
            classes = sorted(d for d in os.listdir(path) if os.path.isdir(os.path.join(path, d)))
            label_map = {c: i + (1 if one_based_label else 0) for i, c in enumerate(classes)}
            for c in classes:
                fs = sorted(f for f in glob.glob(os.path.join(path, c, "*")) if f.lower().endswith(IMG_EXT))
                files += fs
                labels += [label_map[c]] * len(fs)
        elif os.path.isdir(path):
            files = sorted(f for f in glob.glob(os.path.join(path, "**", "*"), recursive=True) if f.lower().endswith(IMG_EXT))
        else:
            files = sorted(glob.glob(path))
        codec = cv2.IMREAD_COLOR if image_codec in (-1, 1) else image_codec

        def load(i):
            f = files[i]
            with open(f, "rb") as fh:
                raw = fh.read()
            mat = cv2.imdecode(np.frombuffer(raw, np.uint8), codec)
            if mat is None:
                return None
            feat = ImageFeature(uri=f, bytes=raw, originalSize=mat.shape)
            if resize_height > 0 and resize_width > 0:
                mat = cv2.resize(mat, (resize_width, resize_height))
            feat["mat"] = mat
            if with_label:
                feat["label"] = np.array([labels[i]], dtype=np.float32)
            return feat

        # decode on a thread pool (cv2 releases the GIL): the reference parallelises over Spark partitions (`min_partitions`)
        workers = max(1, min(16, os.cpu_count() or 1, len(files)))
        if workers > 1 and len(files) >= 8:
            from concurrent.futures import ThreadPoolExecutor

            with ThreadPoolExecutor(workers) as pool:
                feats = [f for f in pool.map(load, range(len(files))) if f is not None]
        else:
            feats = [f for f in map(load, range(len(files))) if f is not None]
        return cls(feats, label_map)

    @classmethod
    def read_sequence_files(cls, path, sc=None, min_partitions=1, class_num=1000):
        """Images with labels from Hadoop SequenceFiles (a file, or a folder of part files): key "name\\nlabel" / "label", value =
        width, height, raw BGR pixels.  Reference: ImageSet.readSequenceFiles, ImageSet.scala:333-352 (records whose label exceeds
        `class_num` are dropped; the features carry the decoded `mat`, `label` [1] and `originalSize`)."""
        from .seqfile import read_image_records

        path = path[len("file://"):] if str(path).startswith("file://") else str(path)
        feats = []
        for name, label, mat in read_image_records(path, class_num):
            feats.append(ImageFeature(uri=name, mat=mat.copy(), label=np.array([label], dtype=np.float32), originalSize=mat.shape))
        return cls(feats)

    readSequenceFiles = read_sequence_files

    @classmethod
    def from_rdds(cls, image_rdd, label_rdd=None, bigdl_type="float"):
        return LocalImageSet(list(image_rdd), None if label_rdd is None else list(label_rdd))

    @classmethod
    def from_image_frame(cls, image_frame, bigdl_type="float"):
        return cls(list(image_frame))

    # ---- ops -----------------------------------------------------------------------------------
    def is_local(self):
        return True

    def is_distributed(self):
        return False

    @property
    def label_map(self):
        return self._label_map

    def transform(self, transformer):
        """Apply now (features are small dict views; the reference is lazy only because RDDs are)."""
        out = []
        for f in self.features:
            g = ImageFeature(f)
            out.append(transformer(g))
        return type(self)(out, self._label_map) if type(self) is ImageSet else ImageSet(out, self._label_map)

    def get_image_features(self):
        return self.features

    def get_image(self, key="floats", to_chw=True):
        return [f.get_image(key, to_chw) for f in self.features]

    def get_label(self):
        return [f.get("label") for f in self.features]

    def get_predict(self, key="predict"):
        return [(f.get("uri"), f.get(key)) for f in self.features]

    def to_image_frame(self, bigdl_type="float"):
        return self.features

    def to_feature_set(self, **kw):
        from ..common import FeatureSet

        x = np.stack([f.get_tensor() for f in self.features])
        labs = [f.get("label") for f in self.features]
        y = None if any(l is None for l in labs) else np.stack([np.asarray(l, dtype=np.float32).reshape(-1) for l in labs])
        if y is not None and y.shape[1] == 1:
            y = y[:, 0]
        return FeatureSet.numpy(x, y, **kw)

    def __len__(self):
        return len(self.features)


class LocalImageSet(ImageSet):
    def __init__(self, image_list=None, label_list=None, jvalue=None, bigdl_type="float"):
        feats = []
        for i, im in enumerate(image_list or []):
            f = ImageFeature(mat=np.asarray(im), originalSize=np.asarray(im).shape)
            if label_list is not None:
                f["label"] = np.asarray(label_list[i], dtype=np.float32).reshape(-1)
            feats.append(f)
        super().__init__(feats)


DistributedImageSet = LocalImageSet
