"""ROI-aware image ops for detection training: the boxes of a feature follow what happens to its image.

Reference: zoo/.../feature/image/RoiTransformer.scala:25-85 (ImageRoiNormalize / ImageRoiHFlip / ImageRoiResize /
ImageRoiProject), RandomSampler.scala:31-45 (the SSD batch samplers: IoU 0.1 ... 0.9 or the whole image),
ImageExpand.scala:30-50, and BigDL's RoiLabel / BatchSampler.  A feature carries `roi` = float32 [G, 5+] rows
(label, xmin, ymin, xmax, ymax[, difficult]); `roiNormalized` says whether boxes are in [0, 1] or pixels.
"""
from __future__ import annotations

import numpy as np

from .imagePreprocessing import ImagePreprocessing

ROI, NORMALIZED = "roi", "roiNormalized"


def _roi(f):
    r = f.get(ROI)
    return None if r is None else np.asarray(r, dtype=np.float32).reshape(-1, max(5, np.asarray(r).shape[-1] if np.asarray(r).ndim > 1 else 5))


class ImageRoiNormalize(ImagePreprocessing):
    """Pixel boxes -> [0, 1] by the current image size."""

    def apply(self, f):
        r = _roi(f)
        if r is not None and not f.get(NORMALIZED, False):
            h, w = np.asarray(f["mat"]).shape[:2]
            r = r.copy()
            r[:, [1, 3]] /= w
            r[:, [2, 4]] /= h
            f[ROI], f[NORMALIZED] = r, True
        return f


class ImageRoiHFlip(ImagePreprocessing):
    """Mirror the boxes (the image itself is flipped by ImageHFlip)."""

    def __init__(self, normalized=True, bigdl_type="float"):
        self.normalized = normalized

    def apply(self, f):
        r = _roi(f)
        if r is not None:
            w = 1.0 if self.normalized else float(np.asarray(f["mat"]).shape[1])
            r = r.copy()
            x1 = w - r[:, 3]
            r[:, 3] = w - r[:, 1]
            r[:, 1] = x1
            f[ROI] = r
        return f


class ImageRoiResize(ImagePreprocessing):
    """Scale pixel boxes after the image was resized from `originalSize` to its current size."""

    def __init__(self, normalized=False, bigdl_type="float"):
        self.normalized = normalized

    def apply(self, f):
        r = _roi(f)
        if r is not None and not self.normalized and "originalSize" in f:
            h, w = np.asarray(f["mat"]).shape[:2]
            oh, ow = f["originalSize"][:2]
            r = r.copy()
            r[:, [1, 3]] *= w / float(ow)
            r[:, [2, 4]] *= h / float(oh)
            f[ROI] = r
        return f


def project_boxes(roi, crop, need_center=True):
    """Express normalised boxes in the coordinates of the normalised crop window (x1, y1, x2, y2); boxes whose centre falls
    outside (when `need_center`) or that do not intersect the window are dropped."""
    if roi is None or len(roi) == 0:
        return roi
    x1, y1, x2, y2 = crop
    cx, cy = (roi[:, 1] + roi[:, 3]) / 2, (roi[:, 2] + roi[:, 4]) / 2
    keep = (roi[:, 3] > x1) & (roi[:, 1] < x2) & (roi[:, 4] > y1) & (roi[:, 2] < y2)
    if need_center:
        keep &= (cx >= x1) & (cx <= x2) & (cy >= y1) & (cy <= y2)
    out = roi[keep].copy()
    w, h = x2 - x1, y2 - y1
    out[:, [1, 3]] = np.clip((out[:, [1, 3]] - x1) / w, 0, 1)
    out[:, [2, 4]] = np.clip((out[:, [2, 4]] - y1) / h, 0, 1)
    return out[(out[:, 3] > out[:, 1]) & (out[:, 4] > out[:, 2])]


class ImageRoiProject(ImagePreprocessing):
    """After a crop / expand recorded in the feature (`cropBox` or `expandBox`, normalised x1,y1,x2,y2 of the new view in the
    old image), move the boxes into the new view."""

    def __init__(self, need_meet_center_constraint=True, bigdl_type="float"):
        self.need_center = need_meet_center_constraint

    def apply(self, f):
        r, box = _roi(f), f.get("cropBox")
        if r is not None and box is not None:
            f[ROI] = project_boxes(r, box, self.need_center)
            f["cropBox"] = None
        return f


class ImageRoiExpand(ImagePreprocessing):
    """ImageExpand that also moves (normalised) boxes into the larger canvas."""

    def __init__(self, means_r=123, means_g=117, means_b=104, min_expand_ratio=1.0, max_expand_ratio=4.0, bigdl_type="float"):
        self.means, self.lo, self.hi = (means_b, means_g, means_r), float(min_expand_ratio), float(max_expand_ratio)

    def apply(self, f):
        mat = np.asarray(f["mat"])
        ratio = np.random.uniform(self.lo, self.hi)
        h, w = mat.shape[:2]
        nh, nw = int(h * ratio), int(w * ratio)
        y, x = np.random.randint(0, nh - h + 1), np.random.randint(0, nw - w + 1)
        out = np.empty((nh, nw, mat.shape[2]), dtype=mat.dtype)
        out[:] = np.asarray(self.means[:mat.shape[2]], dtype=mat.dtype)
        out[y:y + h, x:x + w] = mat
        f["mat"] = out
        r = _roi(f)
        if r is not None and len(r):
            r = r.copy()
            r[:, [1, 3]] = (r[:, [1, 3]] * w + x) / nw
            r[:, [2, 4]] = (r[:, [2, 4]] * h + y) / nh
            f[ROI] = r
        return f


def _iou_one_to_many(box, boxes):
    lt = np.maximum(box[:2], boxes[:, :2])
    rb = np.minimum(box[2:], boxes[:, 2:])
    inter = np.clip(rb - lt, 0, None).prod(1)
    a = (box[2] - box[0]) * (box[3] - box[1])
    b = (boxes[:, 2] - boxes[:, 0]) * (boxes[:, 3] - boxes[:, 1])
    return inter / np.maximum(a + b - inter, 1e-12)


class BatchSampler:
    """One SSD sampler: draw up to `max_trials` windows (scale in [min_scale, max_scale], aspect ratio in [min_ar, max_ar])
    and accept the first whose best IoU with a ground-truth box lies in [min_overlap, max_overlap]."""

    def __init__(self, max_sample=1, max_trials=50, min_scale=0.3, max_scale=1.0, min_aspect_ratio=0.5, max_aspect_ratio=2.0,
                 min_overlap=None, max_overlap=None):
        self.max_sample, self.max_trials = max_sample, max_trials
        self.scale, self.ar = (min_scale, max_scale), (min_aspect_ratio, max_aspect_ratio)
        self.min_overlap, self.max_overlap = min_overlap, max_overlap

    def sample(self, roi, rng=np.random):
        out = []
        for _ in range(self.max_trials):
            if len(out) >= self.max_sample:
                break
            s = rng.uniform(*self.scale)
            lo, hi = max(self.ar[0], s * s), min(self.ar[1], 1.0 / (s * s))
            ar = rng.uniform(lo, hi) if hi > lo else 1.0
            w, h = min(s * np.sqrt(ar), 1.0), min(s / np.sqrt(ar), 1.0)
            x, y = rng.uniform(0, 1 - w), rng.uniform(0, 1 - h)
            box = np.array([x, y, x + w, y + h], dtype=np.float32)
            if self.min_overlap is None and self.max_overlap is None:
                out.append(box)
                continue
            if roi is None or len(roi) == 0:
                continue
            iou = _iou_one_to_many(box, roi[:, 1:5])
            ok = np.ones_like(iou, dtype=bool)
            if self.min_overlap is not None:
                ok &= iou >= self.min_overlap
            if self.max_overlap is not None:
                ok &= iou <= self.max_overlap
            if ok.any():
                out.append(box)
        return out


class ImageRandomSampler(ImagePreprocessing):
    """The SSD crop augmentation: the whole image, or a window whose IoU with some object is at least 0.1 / 0.3 / 0.5 / 0.7 /
    0.9, or any window (max IoU 1.0); one of the accepted windows is chosen at random, the image is cropped to it and the
    boxes are projected (RandomSampler.scala:31-45)."""

    def __init__(self, bigdl_type="float"):
        self.samplers = [BatchSampler(max_trials=1, min_scale=1.0, max_scale=1.0, min_aspect_ratio=1.0, max_aspect_ratio=1.0)] + \
                        [BatchSampler(min_overlap=o) for o in (0.1, 0.3, 0.5, 0.7, 0.9)] + [BatchSampler(max_overlap=1.0)]

    def apply(self, f):
        r = _roi(f)
        if r is not None and not f.get(NORMALIZED, False):
            raise ValueError("ImageRandomSampler expects normalised boxes: apply ImageRoiNormalize first")
        boxes = []
        for s in self.samplers:
            boxes += s.sample(r)
        if not boxes:
            return f
        x1, y1, x2, y2 = boxes[np.random.randint(len(boxes))]
        mat = np.asarray(f["mat"])
        h, w = mat.shape[:2]
        px1, py1, px2, py2 = int(x1 * w), int(y1 * h), max(int(x2 * w), int(x1 * w) + 1), max(int(y2 * h), int(y1 * h) + 1)
        f["mat"] = mat[py1:py2, px1:px2].copy()
        if r is not None:
            f[ROI] = project_boxes(r, (x1, y1, x2, y2), True)
        return f
