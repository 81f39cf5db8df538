"""ImageConfigure / PaddingParam (reference: pyzoo/zoo/models/image/common/image_config.py:28-70, ImageConfigure.scala)."""
from __future__ import annotations


class PaddingParam:
    pass


class ImageConfigure:
    """How an ImageModel turns an ImageSet into predictions: pre-processing chain, post-processing, per-partition
    batch size, label map, optional feature padding."""

    def __init__(self, pre_processor=None, post_processor=None, batch_per_partition=4, label_map=None, feature_padding_param=None, jvalue=None,
                 bigdl_type="float"):
        self.pre_processor, self.post_processor = pre_processor, post_processor
        self.batch_per_partition, self._label_map, self.feature_padding_param = int(batch_per_partition), label_map, feature_padding_param

    def label_map(self):
        return self._label_map
