"""NNImageReader.readImages(path) -> DataFrame with an `image` struct column
(reference: pyzoo/zoo/pipeline/nnframes/nn_image_reader.py, NNImageReader.scala): each row holds
{origin, height, width, nChannels, mode, data} with OpenCV BGR bytes; `resizeH/resizeW` resize while reading."""
from __future__ import annotations

import glob
import os

import numpy as np


class NNImageReader:
    @staticmethod
    def readImages(path, sc=None, minPartitions=1, resizeH=-1, resizeW=-1, image_codec=-1, bigdl_type="float"):
        import cv2
        import pandas as pd

        paths = []
        for p in str(path).split(","):
            p = p.strip()
            if os.path.isdir(p):
                paths += sorted(f for f in glob.glob(os.path.join(p, "**", "*"), recursive=True) if os.path.isfile(f))
            else:
                paths += sorted(glob.glob(p))
        rows = []
        for f in paths:
            img = cv2.imread(f, cv2.IMREAD_UNCHANGED if image_codec == -1 else image_codec)
            if img is None:
                continue
            if resizeH > 0 and resizeW > 0:
                img = cv2.resize(img, (resizeW, resizeH))
            if img.ndim == 2:
                img = img[:, :, None]
            h, w, c = img.shape
            rows.append({"image": {"origin": "file:" + os.path.abspath(f), "height": h, "width": w, "nChannels": c,
                                   "mode": {1: 0, 3: 16, 4: 24}.get(c, 16), "data": np.ascontiguousarray(img)}})
        return pd.DataFrame(rows, columns=["image"])
