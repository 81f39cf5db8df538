"""Detection evaluation: Pascal VOC average precision (11-point 2007 metric or area under the monotone precision envelope)
and the mean over classes.

Reference: zoo/.../objectdetection/common/evaluation/EvalUtil.scala:44-110 (vocAp), :112-205 (per-class true/false positive
marking, difficult boxes ignored, each ground truth matched once), MeanAveragePrecision.scala:33-93 (a ValidationMethod whose
results add up across batches), PascalVocEvaluator.scala:38-125.
"""
from __future__ import annotations

import numpy as np


def voc_ap(rec, prec, use_07_metric=False):
    rec, prec = np.asarray(rec, dtype=np.float64), np.asarray(prec, dtype=np.float64)
    if rec.size == 0:
        return 0.0
    if use_07_metric:
        return float(sum((prec[rec >= t].max() if np.any(rec >= t) else 0.0) for t in np.arange(0.0, 1.1, 0.1)) / 11.0)
    mrec = np.concatenate([[0.0], rec, [1.0]])
    mpre = np.concatenate([[0.0], prec, [0.0]])
    for i in range(mpre.size - 1, 0, -1):
        mpre[i - 1] = max(mpre[i - 1], mpre[i])
    idx = np.where(mrec[1:] != mrec[:-1])[0]
    return float(np.sum((mrec[idx + 1] - mrec[idx]) * mpre[idx + 1]))


def _iou(box, boxes):
    lt = np.maximum(box[:2], boxes[:, :2])
    rb = np.minimum(box[2:], boxes[:, 2:])
    inter = np.clip(rb - lt, 0, None).prod(1)
    return inter / np.maximum((box[2] - box[0]) * (box[3] - box[1]) + (boxes[:, 2] - boxes[:, 0]) * (boxes[:, 3] - boxes[:, 1]) - inter, 1e-12)


class MeanAveragePrecision:
    """Accumulates detections and ground truth image by image; `result()` -> (mAP, per-class AP dict).

    detections: array [N, 6] = (label, score, xmin, ymin, xmax, ymax) (rows with label < 0 are padding);
    ground truth: array [G, 5 or 6] = (label, xmin, ymin, xmax, ymax[, difficult])."""

    def __init__(self, n_classes, use_07_metric=True, iou_threshold=0.5, bg_label=0, class_names=None):
        self.n_classes, self.use07, self.thr, self.bg = int(n_classes), bool(use_07_metric), float(iou_threshold), bg_label
        self.class_names = class_names
        self.records = {c: [] for c in range(self.n_classes) if c != bg_label}    # (score, tp, fp)
        self.npos = {c: 0 for c in self.records}

    def add(self, detections, ground_truth):
        det = np.asarray(detections, dtype=np.float32).reshape(-1, 6)
        det = det[det[:, 0] >= 0]
        gt = np.asarray(ground_truth, dtype=np.float32)
        gt = gt.reshape(-1, gt.shape[-1] if gt.ndim > 1 else 5)
        for c in self.records:
            g = gt[gt[:, 0] == c] if len(gt) else gt
            difficult = g[:, 5] > 0 if (len(g) and g.shape[1] > 5) else np.zeros(len(g), dtype=bool)
            self.npos[c] += int((~difficult).sum())
            used = np.zeros(len(g), dtype=bool)
            d = det[det[:, 0] == c]
            for row in d[np.argsort(-d[:, 1])]:
                tp = fp = 0
                if len(g):
                    iou = _iou(row[2:6], g[:, 1:5])
                    j = int(iou.argmax())
                    if iou[j] > self.thr:
                        if difficult[j]:
                            continue                       # detections of difficult objects count neither way
                        if not used[j]:
                            tp, used[j] = 1, True
                        else:
                            fp = 1
                    else:
                        fp = 1
                else:
                    fp = 1
                self.records[c].append((float(row[1]), tp, fp))
        return self

    def add_batch(self, detections, ground_truths):
        for d, g in zip(detections, ground_truths):
            self.add(d.detach().cpu().numpy() if hasattr(d, "detach") else d, g.detach().cpu().numpy() if hasattr(g, "detach") else g)
        return self

    def __add__(self, other):
        out = MeanAveragePrecision(self.n_classes, self.use07, self.thr, self.bg, self.class_names)
        for c in out.records:
            out.records[c] = self.records[c] + other.records[c]
            out.npos[c] = self.npos[c] + other.npos[c]
        return out

    def average_precisions(self):
        aps = {}
        for c, recs in self.records.items():
            if self.npos[c] == 0:
                continue
            if not recs:
                aps[c] = 0.0
                continue
            arr = np.array(sorted(recs, key=lambda r: -r[0]), dtype=np.float64)
            tp, fp = np.cumsum(arr[:, 1]), np.cumsum(arr[:, 2])
            aps[c] = voc_ap(tp / self.npos[c], tp / np.maximum(tp + fp, 1e-12), self.use07)
        return aps

    def result(self):
        aps = self.average_precisions()
        m = float(np.mean(list(aps.values()))) if aps else 0.0
        names = self.class_names
        return m, {(names[c] if names else c): v for c, v in aps.items()}

    def __repr__(self):
        return f"PascalMeanAveragePrecision is {self.result()[0]:.4f}"


class PascalVocEvaluator(MeanAveragePrecision):
    """`PascalVocEvaluator(imageSet)`: VOC2007 uses the 11-point metric, later years the area metric."""

    def __init__(self, image_set="voc_2007_test", n_classes=21, class_names=None):
        year = "2007" if "2007" in image_set else "2012"
        super().__init__(n_classes, use_07_metric=(year == "2007"), class_names=class_names)
        self.year = year
