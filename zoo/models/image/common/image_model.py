"""ImageModel: a model that predicts over an ImageSet (reference: pyzoo/zoo/models/image/common/image_model.py:27-50,
ImageModel.scala): `predict_image_set(image_set, configure)` runs configure.pre_processor -> the network (batched on
the GPU) -> configure.post_processor, and stores the result under each ImageFeature's `predict` key."""
from __future__ import annotations

import numpy as np
import torch

from ...common.zoo_model import ZooModel


class ImageModel(ZooModel):
    def __init__(self, net=None, config=None, bigdl_type="float"):
        super().__init__()
        self.net = net
        self.config = config

    def forward(self, x):
        return self.net(x)

    def get_config(self):
        return self.config

    def predict_image_set(self, image, configure=None):
        from ....feature.image.imageset import ImageSet

        cfg = configure or self.config
        data = image
        if cfg is not None and cfg.pre_processor is not None:
            data = image.transform(cfg.pre_processor)
        feats = data.get_image_features() if isinstance(data, ImageSet) else list(data)
        tensors = [f.get_tensor() for f in feats]
        dev = next(self.parameters()).device if any(True for _ in self.parameters()) else torch.device("cpu")
        bs = max(1, (cfg.batch_per_partition if cfg else 4)) * 16
        self.eval()
        outs = []
        with torch.no_grad():
            for s in range(0, len(tensors), bs):
                x = torch.from_numpy(np.stack(tensors[s:s + bs])).to(dev)
                o = self.net(x)
                outs.append(o.float().cpu().numpy() if torch.is_tensor(o) else [t.float().cpu().numpy() for t in o])
        preds = np.concatenate(outs) if outs and isinstance(outs[0], np.ndarray) else outs
        for f, p in zip(feats, preds):
            f["predict"] = p
        if cfg is not None and cfg.post_processor is not None:
            for f in feats:
                cfg.post_processor(f)
        return data
