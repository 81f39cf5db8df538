"""ObjectDetector + detection post-processing (reference: pyzoo/zoo/models/image/objectdetection/object_detector.py:30-140;
Scala ObjectDetector.scala, BboxUtil.scala `decodeRois` / `scaleBBox`, Visualizer.scala).

Scope: inference-side pieces -- wrapping a detection network, decoding its `[N, 6]` output rows
(label, score, x1, y1, x2, y2), scaling boxes back to the original image, non-maximum suppression, and drawing.
SSD training (`MultiBoxLoss`, prior boxes, mAP evaluators, 2.5 kLoC in the reference) is not implemented."""
from __future__ import annotations

import numpy as np
import torch

from ..common.image_config import ImageConfigure
from ..common.image_model import ImageModel

PASCAL_CLASSES = ["__background__", "aeroplane", "bicycle", "bird", "boat", "bottle", "bus", "car", "cat", "chair", "cow", "diningtable", "dog",
                  "horse", "motorbike", "person", "pottedplant", "sheep", "sofa", "train", "tvmonitor"]


def read_pascal_label_map():
    return dict(enumerate(PASCAL_CLASSES))


def read_coco_label_map():
    return {i: f"coco_{i}" for i in range(81)}


class ImInfo:
    """Stores (height, width, scaleH, scaleW) of the pre-processed image under `ImInfo`."""

    def __call__(self, f):
        h, w = f["mat"].shape[:2]
        oh, ow = f.get("originalSize", (h, w, 3))[:2]
        f["ImInfo"] = np.array([h, w, h / oh, w / ow], dtype=np.float32)
        return f


class DecodeOutput:
    """Flat detection output `[1 + 6k]` (count, then rows) or `[k, 6]` -> `[k, 6]` rows (label, score, x1, y1, x2, y2)."""

    def __call__(self, f):
        p = np.asarray(f["predict"], dtype=np.float32)
        if p.ndim == 1:
            n = int(p[0])
            p = p[1:1 + 6 * n].reshape(n, 6)
        f["predict"] = p.reshape(-1, 6)
        return f


class ScaleDetection:
    """Normalised (0..1) boxes -> pixel coordinates of the ORIGINAL image, clipped to its bounds."""

    def __call__(self, f):
        p = np.asarray(f["predict"], dtype=np.float32).reshape(-1, 6).copy()
        h, w = f.get("originalSize", f["mat"].shape)[:2]
        p[:, [2, 4]] = np.clip(p[:, [2, 4]] * w, 0, w - 1)
        p[:, [3, 5]] = np.clip(p[:, [3, 5]] * h, 0, h - 1)
        f["predict"] = p
        return f


def nms(boxes, scores, iou_threshold=0.45, top_k=200):
    """Greedy non-maximum suppression (BboxUtil.scala `nms`); returns kept indices, best first."""
    boxes, scores = np.asarray(boxes, dtype=np.float32), np.asarray(scores, dtype=np.float32)
    order = np.argsort(-scores)[:top_k]
    keep = []
    area = (boxes[:, 2] - boxes[:, 0]).clip(0) * (boxes[:, 3] - boxes[:, 1]).clip(0)
    while len(order):
        i = order[0]
        keep.append(int(i))
        rest = order[1:]
        xx1, yy1 = np.maximum(boxes[i, 0], boxes[rest, 0]), np.maximum(boxes[i, 1], boxes[rest, 1])
        xx2, yy2 = np.minimum(boxes[i, 2], boxes[rest, 2]), np.minimum(boxes[i, 3], boxes[rest, 3])
        inter = (xx2 - xx1).clip(0) * (yy2 - yy1).clip(0)
        iou = inter / np.maximum(area[i] + area[rest] - inter, 1e-12)
        order = rest[iou <= iou_threshold]
    return keep


class Visualizer:
    """Draws boxes with score >= thresh onto the image; result bytes (png/jpg) under `out_key`."""

    def __init__(self, label_map, thresh=0.3, encoding="png", bigdl_type="float", out_key="visualized"):
        self.label_map, self.thresh, self.encoding, self.out_key = label_map, float(thresh), encoding, out_key

    def __call__(self, image_set):
        import cv2

        feats = image_set.get_image_features() if hasattr(image_set, "get_image_features") else [image_set]
        for f in feats:
            img = np.ascontiguousarray(f.get("originalMat", f["mat"])).astype(np.uint8).copy()
            for lab, score, x1, y1, x2, y2 in np.asarray(f["predict"]).reshape(-1, 6):
                if score < self.thresh:
                    continue
                cv2.rectangle(img, (int(x1), int(y1)), (int(x2), int(y2)), (0, 0, 255), 2)
                name = self.label_map.get(int(lab), str(int(lab))) if self.label_map else str(int(lab))
                cv2.putText(img, f"{name}: {score:.2f}", (int(x1), max(int(y1) - 4, 10)), cv2.FONT_HERSHEY_SIMPLEX, 0.5, (0, 0, 255), 1)
            ok, buf = cv2.imencode("." + self.encoding, img)
            f[self.out_key] = buf.tobytes() if ok else None
        return image_set


class ObjectDetector(ImageModel):
    def __init__(self, net=None, config=None, label_map=None, bigdl_type="float"):
        if config is None:
            config = ImageConfigure(label_map=label_map or read_pascal_label_map(), batch_per_partition=2)
        super().__init__(net, config)

    @staticmethod
    def load_model(path, weight_path=None, bigdl_type="float"):
        try:
            net = torch.jit.load(path, map_location="cpu")
        except Exception:
            net = torch.load(path, map_location="cpu", weights_only=False)
        if weight_path is not None:
            net.load_state_dict(torch.load(weight_path, map_location="cpu"))
        return ObjectDetector(net)
