"""ImageClassifier (reference: pyzoo/zoo/models/image/imageclassification/image_classification.py:27-80,
ImageClassifier.scala, ImageClassificationConfig.scala).  The reference loads published BigDL models by name; here a
classifier wraps any torch network (e.g. torchvision `resnet50`) with the reference's per-model pre-processing
(resize 256 -> centre crop 224 -> channel normalise) and `LabelOutput` post-processing (top classes + probabilities)."""
from __future__ import annotations

import numpy as np
import torch

from ..common.image_config import ImageConfigure
from ..common.image_model import ImageModel


def read_imagenet_label_map():
    """index -> class name.  The label file ships with the reference jar; offline a numeric map is returned."""
    return {i: f"class_{i}" for i in range(1000)}


class LabelOutput:
    """Post-processor: predict vector -> `clses` (labels sorted by probability) and `probs` keys."""

    def __init__(self, label_map, clses="clses", probs="probs", bigdl_type="float"):
        self.label_map, self.clses, self.probs = label_map, clses, probs

    def __call__(self, feature):
        p = np.asarray(feature["predict"]).reshape(-1)
        order = np.argsort(-p)
        feature[self.clses] = [self.label_map.get(int(i), str(int(i))) if self.label_map else int(i) for i in order]
        feature[self.probs] = p[order]
        return feature


class ImageClassifier(ImageModel):
    def __init__(self, net=None, config=None, bigdl_type="float"):
        if config is None:
            from ....feature.image import ChainedPreprocessing, ImageCenterCrop, ImageChannelNormalize, ImageMatToTensor, ImageResize

            pre = ChainedPreprocessing([ImageResize(256, 256), ImageCenterCrop(224, 224),
                                        ImageChannelNormalize(123.68, 116.779, 103.939, 58.393, 57.12, 57.375), ImageMatToTensor()])
            config = ImageConfigure(pre_processor=pre, post_processor=LabelOutput(read_imagenet_label_map()), batch_per_partition=4,
                                    label_map=read_imagenet_label_map())
        super().__init__(net, config)

    @staticmethod
    def load_model(path, weight_path=None, bigdl_type="float"):
        """A TorchScript archive / pickled torch module (the reference's BigDL `.model` files cannot be read here)."""
        try:
            net = torch.jit.load(path, map_location="cpu")
        except Exception:
            net = torch.load(path, map_location="cpu", weights_only=False)
        if weight_path is not None:
            net.load_state_dict(torch.load(weight_path, map_location="cpu"))
        return ImageClassifier(net)
