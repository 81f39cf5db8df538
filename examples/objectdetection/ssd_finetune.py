"""SSD fine-tuning on a synthetic detection set: ROI augmentation chain -> SSD-300 heads -> MultiBoxLoss -> VOC mAP.
(reference example: zoo/.../examples/objectdetection/finetune/ssd/Train.scala)

    python examples/objectdetection/ssd_finetune.py [--steps 30] [--device cuda]

The data are coloured rectangles on noise (class = colour); the point is the pipeline, which is the reference's:
RoiNormalize -> ColorJitter -> Expand -> RandomSampler -> Resize(300) -> HFlip(0.5) -> batch -> loss."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch

from zoo.feature.image.imagePreprocessing import ImageColorJitter, ImageHFlip, ImageResize
from zoo.feature.image.imageset import ImageFeature
from zoo.feature.image.roi import ImageRandomSampler, ImageRoiExpand, ImageRoiHFlip, ImageRoiNormalize
from zoo.models.image.objectdetection import MeanAveragePrecision, MultiBoxLoss, SSDVGG, roi_batch

COLORS = {1: (255, 40, 40), 2: (40, 255, 40), 3: (40, 40, 255)}


def make_feature(rng):
    mat = rng.randint(0, 60, (240, 320, 3)).astype(np.uint8)
    rows = []
    for _ in range(rng.randint(1, 3)):
        lab = int(rng.randint(1, 4))
        x1, y1 = rng.randint(0, 200), rng.randint(0, 140)
        w, h = rng.randint(60, 110), rng.randint(50, 90)
        mat[y1:y1 + h, x1:x1 + w] = COLORS[lab]
        rows.append([lab, x1, y1, x1 + w, y1 + h])
    return ImageFeature(mat=mat, roi=np.asarray(rows, np.float32), originalSize=mat.shape, roiNormalized=False)


def augment(f, train):
    f = ImageRoiNormalize().apply(f)
    if train:
        f = ImageColorJitter().apply(f)
        f = ImageRoiExpand(max_expand_ratio=2.0).apply(f)
        f = ImageRandomSampler().apply(f)
    f = ImageResize(300, 300).apply(f)
    if train and np.random.rand() < 0.5:
        f = ImageRoiHFlip().apply(ImageHFlip().apply(f))
    return f


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    a = ap.parse_args()
    rng = np.random.RandomState(0)
    np.random.seed(0)
    torch.manual_seed(0)
    dev = torch.device(a.device)
    net = SSDVGG(class_num=4, resolution=300, conf_thresh=0.3).to(dev)
    crit = MultiBoxLoss(n_classes=4, priors=net.priors)
    opt = torch.optim.SGD(net.parameters(), lr=1e-3, momentum=0.9, weight_decay=5e-4)
    net.train()
    for step in range(a.steps):
        x, t = roi_batch([augment(make_feature(rng), True) for _ in range(a.batch)])
        loss = crit(net(x.to(dev)), t.to(dev))
        opt.zero_grad()
        loss.backward()
        opt.step()
        if step % 5 == 0 or step == a.steps - 1:
            print(f"step {step:3d}  multibox loss {float(loss):.4f}")
    net.eval()
    m = MeanAveragePrecision(4, use_07_metric=True)
    with torch.no_grad():
        for _ in range(4):
            feats = [augment(make_feature(rng), False) for _ in range(a.batch)]
            x, t = roi_batch(feats)
            det = net(x.to(dev)).cpu()
            m.add_batch(det, [t[t[:, 0] == i][:, 1:] for i in range(a.batch)])
    print(m)


if __name__ == "__main__":
    main()
