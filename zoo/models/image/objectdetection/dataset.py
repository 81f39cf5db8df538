"""Detection datasets: Pascal VOC (xml annotations) and COCO (json) readers producing ROI image features.

Reference: zoo/.../objectdetection/common/dataset/PascalVoc.scala:33-114 (devkit layout `VOC<year>/{ImageSets/Main,
Annotations,JPEGImages}`, 20 classes + background, `difficult` flag), Coco.scala:40-188 (80 categories mapped to contiguous
labels, crowd boxes as difficult), Imdb.scala:30-95 (`Imdb.getImdb(name, devkitPath)`), CustomizedDataSet.scala.
"""
from __future__ import annotations

import json
import os
import xml.etree.ElementTree as ET

import numpy as np

PASCAL_CLASSES = ("__background__", "aeroplane", "bicycle", "bird", "boat", "bottle", "bus", "car", "cat", "chair", "cow",
                  "diningtable", "dog", "horse", "motorbike", "person", "pottedplant", "sheep", "sofa", "train", "tvmonitor")


class Imdb:
    """Base: `roidb()` -> list of dicts {uri, path, roi [G, 6] = (label, x1, y1, x2, y2, difficult) in pixels}."""
    classes = ()

    def roidb(self):  # pragma: no cover - interface
        raise NotImplementedError

    def to_image_set(self, read_image=True):
        from ....feature.image.imageset import ImageFeature, ImageSet

        feats = []
        for r in self.roidb():
            f = ImageFeature(uri=r["uri"], roi=r["roi"], roiNormalized=False)
            if read_image:
                import cv2

                mat = cv2.imread(r["path"], cv2.IMREAD_COLOR)
                if mat is None:
                    raise ValueError(f"cannot read {r['path']}")
                f["mat"], f["originalSize"] = mat, mat.shape
            else:
                f["path"] = r["path"]
            feats.append(f)
        return ImageSet(feats, label_map={n: i for i, n in enumerate(self.classes)})

    @staticmethod
    def get_imdb(name, devkit_path):
        """`voc_2007_trainval`, `voc_2012_test`, `coco_train2014`, ... (Imdb.scala:77-94)."""
        parts = name.split("_")
        if parts[0] == "voc":
            return PascalVoc(parts[1], parts[2], devkit_path)
        if parts[0] == "coco":
            return Coco(parts[1], devkit_path)
        raise ValueError(f"unknown imdb {name}")


class PascalVoc(Imdb):
    classes = PASCAL_CLASSES

    def __init__(self, year="2007", image_set="trainval", devkit_path="."):
        self.year, self.image_set, self.root = str(year), image_set, os.path.join(devkit_path, f"VOC{year}")
        self.cls_to_ind = {c: i for i, c in enumerate(self.classes)}

    def image_index(self):
        with open(os.path.join(self.root, "ImageSets", "Main", self.image_set + ".txt")) as fh:
            return [l.split()[0] for l in fh if l.strip()]

    def load_annotation(self, index):
        tree = ET.parse(os.path.join(self.root, "Annotations", index + ".xml"))
        rows = []
        for obj in tree.findall("object"):
            name = obj.find("name").text.strip().lower()
            if name not in self.cls_to_ind:
                continue
            bb = obj.find("bndbox")
            d = obj.find("difficult")
            # VOC pixel indices are 1-based
            rows.append([self.cls_to_ind[name]] + [float(bb.find(k).text) - 1 for k in ("xmin", "ymin", "xmax", "ymax")] +
                        [float(d.text) if d is not None else 0.0])
        return np.asarray(rows, dtype=np.float32).reshape(-1, 6)

    def roidb(self):
        return [{"uri": i, "path": os.path.join(self.root, "JPEGImages", i + ".jpg"), "roi": self.load_annotation(i)} for i in self.image_index()]


class Coco(Imdb):
    def __init__(self, image_set="train2014", devkit_path="."):
        self.image_set, self.root = image_set, devkit_path
        self._ann = None

    def _load(self):
        if self._ann is None:
            with open(os.path.join(self.root, "annotations", f"instances_{self.image_set}.json")) as fh:
                self._ann = json.load(fh)
            cats = sorted(self._ann["categories"], key=lambda c: c["id"])
            self.classes = ("__background__",) + tuple(c["name"] for c in cats)
            self.cat_to_label = {c["id"]: i + 1 for i, c in enumerate(cats)}
        return self._ann

    def roidb(self):
        ann = self._load()
        by_img = {}
        for a in ann["annotations"]:
            x, y, w, h = a["bbox"]
            if w <= 0 or h <= 0:
                continue
            by_img.setdefault(a["image_id"], []).append([self.cat_to_label[a["category_id"]], x, y, x + w - 1, y + h - 1, float(a.get("iscrowd", 0))])
        out = []
        for im in ann["images"]:
            roi = np.asarray(by_img.get(im["id"], []), dtype=np.float32).reshape(-1, 6)
            out.append({"uri": str(im["id"]), "path": os.path.join(self.root, self.image_set, im["file_name"]), "roi": roi})
        return out


class CustomizedDataSet(Imdb):
    """A devkit in Pascal layout with its own class list (`classes.txt` under the root, one name per line)."""

    def __init__(self, name, image_set, devkit_path):
        self.inner = PascalVoc.__new__(PascalVoc)
        self.inner.year, self.inner.image_set, self.inner.root = name, image_set, os.path.join(devkit_path, name)
        with open(os.path.join(self.inner.root, "classes.txt")) as fh:
            self.classes = ("__background__",) + tuple(l.strip() for l in fh if l.strip())
        self.inner.classes = self.classes
        self.inner.cls_to_ind = {c: i for i, c in enumerate(self.classes)}

    def roidb(self):
        return self.inner.roidb()


def roi_batch(features, resolution=300, mean_bgr=(104.0, 117.0, 123.0)):
    """Image features (already resized to resolution x resolution, normalised boxes) -> (images [B, 3, H, W] float32 RGB-mean
    subtracted BGR planes, target [N, 7] = (image index, label, x1, y1, x2, y2, difficult)): the SSDMiniBatch layout
    (ssd/RoiImageToSSDBatch.scala:40-140)."""
    import torch

    imgs, rows = [], []
    for i, f in enumerate(features):
        m = np.asarray(f["mat"], dtype=np.float32) - np.asarray(mean_bgr, dtype=np.float32)
        imgs.append(torch.from_numpy(np.ascontiguousarray(m.transpose(2, 0, 1))))
        r = np.asarray(f.get("roi", np.zeros((0, 6))), dtype=np.float32)
        if r.size:
            r = r.reshape(-1, r.shape[-1])
            if r.shape[1] == 5:
                r = np.concatenate([r, np.zeros((len(r), 1), np.float32)], 1)
            rows.append(np.concatenate([np.full((len(r), 1), i, np.float32), r[:, :6]], 1))
    target = torch.from_numpy(np.concatenate(rows, 0)) if rows else torch.zeros(0, 7)
    return torch.stack(imgs), target


# ---- ROI records in Hadoop SequenceFiles ----------------------------------------------------------------------------------------
# Reference: RoiByteImageToSeq.scala:36-104 (writer), RoiImageSeqGenerator.scala (command line), Imdb.loadRoiSeqFiles +
# RoiRecordToFeature (reader), ssd/SSDDataSet.scala:31-66 (the training / validation chains over those files).
# Record: key = image path; value = int image-bytes length, int class length (4 bytes per box), the ENCODED image file, then the
# class tensor [2, N] (labels, then difficult flags) and the box tensor [N, 4], big-endian floats.

def encode_roi_record(image_bytes, roi=None):
    import struct

    roi = None if roi is None else np.asarray(roi, dtype=np.float32).reshape(-1, 6)
    n = 0 if roi is None else len(roi)
    out = struct.pack(">ii", len(image_bytes), n * 4) + bytes(image_bytes)
    if n:
        out += roi[:, 0].astype(">f4").tobytes() + roi[:, 5].astype(">f4").tobytes() + roi[:, 1:5].astype(">f4").tobytes()
    return out


def decode_roi_record(value):
    """-> (encoded image bytes, roi [N, 6] = (label, x1, y1, x2, y2, difficult) or None for an unlabelled record)."""
    import struct

    img_len, class_len = struct.unpack_from(">ii", value, 0)
    img = bytes(value[8:8 + img_len])
    n = class_len // 4
    if n == 0:
        return img, None
    pos = 8 + img_len
    cls = np.frombuffer(value, ">f4", 2 * n, pos).astype(np.float32).reshape(2, n)
    box = np.frombuffer(value, ">f4", 4 * n, pos + 8 * n).astype(np.float32).reshape(n, 4)
    return img, np.concatenate([cls[0][:, None], box, cls[1][:, None]], 1)


class RoiByteImageToSeq:
    """roidb entries -> SequenceFiles `<base>_<k>.seq` of `block_size` records each; returns the file names."""

    def __init__(self, block_size, base_file_name):
        self.block_size, self.base = int(block_size), str(base_file_name)

    def __call__(self, roidb):
        from ....feature.image.seqfile import write_sequence_file

        files, entries = [], list(roidb)
        for k in range(0, max(len(entries), 1), self.block_size):
            chunk = entries[k:k + self.block_size]
            if not chunk:
                break
            recs = []
            for r in chunk:
                with open(r["path"], "rb") as fh:
                    recs.append((str(r["path"]).encode(), encode_roi_record(fh.read(), r.get("roi"))))
            name = f"{self.base}_{k // self.block_size}.seq"
            write_sequence_file(name, recs)
            files.append(name)
        return files


def load_roi_seq_files(folder):
    """Every record of the `.seq` files under `folder` as an ImageFeature with `bytes`, `uri` and (when labelled) `roi`
    (Imdb.loadRoiSeqFiles + RoiRecordToFeature(convertLabel = true))."""
    import glob

    from ....feature.image.imageset import ImageFeature
    from ....feature.image.seqfile import read_sequence_file

    feats = []
    for p in sorted(glob.glob(os.path.join(folder, "*.seq"))):
        for key, value in read_sequence_file(p):
            img, roi = decode_roi_record(value)
            f = ImageFeature(uri=key.decode(), bytes=img)
            if roi is not None:
                f["roi"], f["roiNormalized"] = roi, False
            feats.append(f)
    return feats


class RoiImageSeqGenerator:
    """`python -m zoo.models.image.objectdetection.dataset -f <devkit or image folder> -o <out> [-i voc_2007_trainval] [-b 12800] [-p 1]`"""

    @staticmethod
    def generate(folder, output, image_set=None, block_size=12800, parallel=1):
        os.makedirs(output, exist_ok=True)
        if image_set is None:                               # unlabelled images
            roidb = [{"uri": f, "path": os.path.join(folder, f)} for f in sorted(os.listdir(folder))
                     if f.lower().endswith((".jpg", ".jpeg", ".png", ".bmp"))]
        else:
            roidb = Imdb.get_imdb(image_set, folder).roidb()
        parallel = max(1, min(int(parallel), len(roidb) or 1))
        shares = [roidb[t::parallel] for t in range(parallel)]
        files = []
        from concurrent.futures import ThreadPoolExecutor

        with ThreadPoolExecutor(parallel) as pool:
            for fs in pool.map(lambda ts: RoiByteImageToSeq(block_size, os.path.join(output, f"{len(roidb)}-seq-{ts[0]}"))(ts[1]), enumerate(shares)):
                files += fs
        return files

    @staticmethod
    def main(argv=None):
        import argparse

        ap = argparse.ArgumentParser(description="Pascal VOC / COCO / image folder -> ROI SequenceFiles")
        ap.add_argument("-f", "--folder", required=True, help="image folder, or the devkit path for labelled data")
        ap.add_argument("-o", "--output", required=True)
        ap.add_argument("-p", "--parallel", type=int, default=1)
        ap.add_argument("-b", "--blockSize", type=int, default=12800)
        ap.add_argument("-i", "--imageSet", default=None, help="e.g. voc_2007_test; omit for unlabelled images")
        a = ap.parse_args(argv)
        for f in RoiImageSeqGenerator.generate(a.folder, a.output, a.imageSet, a.blockSize, a.parallel):
            print("Generated file", f)


class SSDDataSet:
    """The reference's SSD input chains over ROI SequenceFiles (ssd/SSDDataSet.scala:31-66); each yields batches
    (images [B, 3, R, R], target [N, 7]) in the SSDMiniBatch layout."""

    @staticmethod
    def _batches(feats, chain, batch_size, resolution, shuffle):
        order = np.random.permutation(len(feats)) if shuffle else np.arange(len(feats))
        for k in range(0, len(order), batch_size):
            yield roi_batch([chain(dict(feats[i])) for i in order[k:k + batch_size]], resolution)

    @staticmethod
    def load_ssd_train_set(folder, sc=None, resolution=300, batch_size=32, par_num=None):
        from ....feature.image.imagePreprocessing import (ImageBytesToMat, ImageColorJitter, ImageHFlip, ImageRandomPreprocessing, ImageResize)
        from ....feature.image.roi import ImageRandomSampler, ImageRoiExpand, ImageRoiHFlip, ImageRoiNormalize
        from ....feature.common import ChainedPreprocessing

        feats = load_roi_seq_files(folder)
        flip = ChainedPreprocessing([ImageHFlip(), ImageRoiHFlip()])
        steps = [ImageBytesToMat(), ImageRoiNormalize(), ImageColorJitter(), ImageRandomPreprocessing(ImageRoiExpand(), 0.5), ImageRandomSampler(),
                 ImageResize(resolution, resolution, -1), ImageRandomPreprocessing(flip, 0.5)]

        def chain(f):
            from ....feature.image.imageset import ImageFeature

            f = ImageFeature(**f)
            for s in steps:
                f = s(f)
            return f

        return lambda: SSDDataSet._batches(feats, chain, batch_size, resolution, True)

    @staticmethod
    def load_ssd_val_set(folder, sc=None, resolution=300, batch_size=32, par_num=None):
        from ....feature.image.imagePreprocessing import ImageBytesToMat, ImageResize
        from ....feature.image.roi import ImageRoiNormalize

        feats = load_roi_seq_files(folder)
        steps = [ImageBytesToMat(), ImageRoiNormalize(), ImageResize(resolution, resolution)]

        def chain(f):
            from ....feature.image.imageset import ImageFeature

            f = ImageFeature(**f)
            for s in steps:
                f = s(f)
            return f

        return lambda: SSDDataSet._batches(feats, chain, batch_size, resolution, False)


if __name__ == "__main__":
    RoiImageSeqGenerator.main()
