"""SSD stack on CPU: priors, box coding, matching, MultiBox loss, detection output, VOC mAP, ROI transforms, VOC reader."""
import os

import numpy as np
import pytest
import torch

from zoo.models.image.objectdetection import (MeanAveragePrecision, MultiBoxLoss, PascalVoc, SSDVGG, DetectionOutput, decode_boxes,
                                              encode_boxes, jaccard, match_priors, nms, prior_boxes, roi_batch, ssd_specs, voc_ap)


def test_prior_counts_match_the_published_ssd_configs():
    assert prior_boxes(ssd_specs(300), 300).shape == (8732, 4)
    assert prior_boxes(ssd_specs(512), 512).shape == (24564, 4)
    p = prior_boxes(ssd_specs(300), 300)
    # first prior of the 38x38 map: centre (0.5*8/300), size 30/300
    torch.testing.assert_close(p[0], torch.tensor([4 - 15, 4 - 15, 4 + 15, 4 + 15]) / 300.0)
    with pytest.raises(ValueError):
        ssd_specs(400)
    with pytest.raises(ValueError):
        ssd_specs(300, sizes=[1, 2, 3])


def test_encode_decode_roundtrip_and_iou():
    torch.manual_seed(0)
    pri = prior_boxes(ssd_specs(300), 300)[1000:1100]
    gt = (pri + 0.03 * torch.randn_like(pri)).clamp(0, 1)
    gt = torch.cat([torch.min(gt[:, :2], gt[:, 2:]) - 0.01, torch.max(gt[:, :2], gt[:, 2:]) + 0.01], 1)
    torch.testing.assert_close(decode_boxes(encode_boxes(gt, pri), pri), gt, atol=1e-5, rtol=1e-5)
    a = torch.tensor([[0.0, 0.0, 0.5, 0.5]])
    b = torch.tensor([[0.25, 0.25, 0.75, 0.75], [0.0, 0.0, 0.5, 0.5], [0.6, 0.6, 0.9, 0.9]])
    torch.testing.assert_close(jaccard(a, b), torch.tensor([[0.0625 / 0.4375, 1.0, 0.0]]))


def test_nms_keeps_best_of_each_cluster():
    boxes = torch.tensor([[0.1, 0.1, 0.4, 0.4], [0.12, 0.1, 0.42, 0.4], [0.6, 0.6, 0.9, 0.9], [0.61, 0.6, 0.9, 0.91]])
    scores = torch.tensor([0.8, 0.9, 0.7, 0.3])
    assert nms(boxes, scores, 0.45).tolist() == [1, 2]


def test_matching_bipartite_then_threshold():
    pri = torch.tensor([[0.0, 0.0, 0.2, 0.2], [0.1, 0.1, 0.3, 0.3], [0.5, 0.5, 0.9, 0.9], [0.52, 0.5, 0.92, 0.9]])
    gt = torch.tensor([[0.05, 0.05, 0.25, 0.25], [0.5, 0.5, 0.9, 0.9]])
    m, iou = match_priors(gt, pri, 0.5)
    assert m[2] == 1 and m[3] == 1            # exact match + a high-overlap neighbour over the threshold
    assert (m[:2] >= 0).sum() == 1             # gt 0 claims exactly one of the two low-overlap priors (bipartite step)
    m2, _ = match_priors(torch.zeros(0, 4), pri)
    assert (m2 == -1).all()


def test_multibox_loss_trains_a_tiny_head_and_detection_output_finds_the_box():
    torch.manual_seed(1)
    pri = prior_boxes(ssd_specs(300), 300)[-200:]                 # the coarse maps only, to keep the test small
    P, C = pri.shape[0], 3
    loc = torch.zeros(2, P, 4, requires_grad=True)
    conf = torch.zeros(2, P, C, requires_grad=True)
    target = torch.tensor([[0, 1, 0.1, 0.1, 0.6, 0.7, 0], [1, 2, 0.3, 0.2, 0.9, 0.8, 0]])
    crit = MultiBoxLoss(n_classes=C, priors=pri)
    opt = torch.optim.Adam([loc, conf], lr=0.1)
    first = None
    for _ in range(60):
        opt.zero_grad()
        loss = crit((loc, conf), target)
        loss.backward()
        opt.step()
        first = first if first is not None else float(loss)
    assert float(loss) < 0.2 * first
    det = DetectionOutput(C, conf_thresh=0.3)(loc.detach(), conf.detach(), pri)
    assert det.shape == (2, 200, 6)
    for b, (lab, box) in enumerate([(1, [0.1, 0.1, 0.6, 0.7]), (2, [0.3, 0.2, 0.9, 0.8])]):
        top = det[b, 0]
        assert int(top[0]) == lab and float(top[1]) > 0.5
        assert float(jaccard(top[None, 2:], torch.tensor([box]))[0, 0]) > 0.7
    # list-of-tensors target form gives the same value
    l2 = crit((loc, conf), [target[:1, 1:], target[1:, 1:]])
    torch.testing.assert_close(l2, crit((loc, conf), target))


def test_ssdvgg_shapes_train_and_eval():
    net = SSDVGG(class_num=4, resolution=300)
    x = torch.randn(1, 3, 300, 300)
    net.train()
    loc, conf = net(x)
    assert loc.shape == (1, 8732, 4) and conf.shape == (1, 8732, 4)
    MultiBoxLoss(n_classes=4, priors=net.priors)((loc, conf), torch.tensor([[0, 1, 0.2, 0.2, 0.7, 0.7, 0.0]])).backward()
    assert net.loc[0].weight.grad is not None
    net.eval()
    with torch.no_grad():
        d = SSDVGG(class_num=4, resolution=300, conf_thresh=0.9).eval()(x)
    assert d.shape == (1, 200, 6)


def test_voc_ap_and_mean_average_precision():
    assert voc_ap([0.5, 1.0], [1.0, 1.0], True) == pytest.approx(1.0)
    assert voc_ap([0.5, 1.0], [1.0, 0.5], False) == pytest.approx(0.75)
    m = MeanAveragePrecision(3, use_07_metric=False)
    gt = np.array([[1, 0.1, 0.1, 0.5, 0.5, 0], [2, 0.6, 0.6, 0.9, 0.9, 0], [1, 0.7, 0.1, 0.9, 0.3, 1]])      # the third is difficult
    det = np.array([[1, 0.9, 0.1, 0.1, 0.5, 0.5], [1, 0.8, 0.12, 0.1, 0.5, 0.5], [2, 0.7, 0.0, 0.0, 0.2, 0.2], [1, 0.6, 0.7, 0.1, 0.9, 0.3],
                    [-1, -1, -1, -1, -1, -1]])
    m.add(det, gt)
    mp, aps = m.result()
    assert aps[1] == pytest.approx(1.0)      # one TP first, the duplicate is a FP after full recall, the difficult hit is ignored
    assert aps[2] == pytest.approx(0.0)
    assert mp == pytest.approx(0.5)
    both = m + m
    assert both.npos[1] == 2


def test_roi_transforms_and_sampler():
    from zoo.feature.image.imageset import ImageFeature
    from zoo.feature.image.roi import ImageRandomSampler, ImageRoiExpand, ImageRoiHFlip, ImageRoiNormalize, ImageRoiResize, project_boxes

    np.random.seed(0)
    f = ImageFeature(mat=np.zeros((100, 200, 3), np.uint8), roi=np.array([[1, 20, 10, 120, 60]], np.float32), originalSize=(100, 200, 3))
    f = ImageRoiNormalize().apply(f)
    np.testing.assert_allclose(f["roi"][0, 1:], [0.1, 0.1, 0.6, 0.6])
    f = ImageRoiHFlip().apply(f)
    np.testing.assert_allclose(f["roi"][0, 1:], [0.4, 0.1, 0.9, 0.6], atol=1e-6)
    out = project_boxes(f["roi"], (0.5, 0.0, 1.0, 1.0))
    np.testing.assert_allclose(out[0, 1:], [0.0, 0.1, 0.8, 0.6], atol=1e-6)
    assert len(project_boxes(f["roi"], (0.0, 0.7, 1.0, 1.0))) == 0
    g = ImageRoiExpand(min_expand_ratio=2.0, max_expand_ratio=2.0).apply(ImageFeature(mat=np.ones((10, 10, 3), np.uint8),
                                                                                    roi=np.array([[1, 0, 0, 1, 1]], np.float32), roiNormalized=True))
    assert g["mat"].shape == (20, 20, 3)
    b = g["roi"][0]
    assert b[3] - b[1] == pytest.approx(0.5) and b[4] - b[2] == pytest.approx(0.5)
    h = ImageRandomSampler().apply(ImageFeature(mat=np.zeros((120, 160, 3), np.uint8), roi=np.array([[1, 0.2, 0.2, 0.8, 0.8]], np.float32),
                                                roiNormalized=True))
    assert h["mat"].shape[0] <= 120 and h["mat"].shape[1] <= 160 and (len(h["roi"]) == 0 or h["roi"][:, 1:].max() <= 1.0)
    px = ImageFeature(mat=np.zeros((50, 100, 3), np.uint8), roi=np.array([[1, 20, 10, 120, 60]], np.float32), originalSize=(100, 200, 3))
    np.testing.assert_allclose(ImageRoiResize().apply(px)["roi"][0, 1:], [10, 5, 60, 30])


def test_pascal_voc_reader_and_roi_batch(tmp_path):
    import cv2

    root = tmp_path / "VOC2007"
    for d in ("ImageSets/Main", "Annotations", "JPEGImages"):
        os.makedirs(root / d)
    (root / "ImageSets/Main/trainval.txt").write_text("000001\n000002 1\n")
    for idx, (name, box) in {"000001": ("dog", (11, 21, 101, 151)), "000002": ("person", (5, 6, 50, 60))}.items():
        cv2.imwrite(str(root / "JPEGImages" / f"{idx}.jpg"), np.full((200, 300, 3), 128, np.uint8))
        (root / "Annotations" / f"{idx}.xml").write_text(
            f"<annotation><object><name>{name}</name><difficult>0</difficult><bndbox><xmin>{box[0]}</xmin><ymin>{box[1]}</ymin>"
            f"<xmax>{box[2]}</xmax><ymax>{box[3]}</ymax></bndbox></object><object><name>unicorn</name><bndbox><xmin>1</xmin><ymin>1</ymin>"
            f"<xmax>2</xmax><ymax>2</ymax></bndbox></object></annotation>")
    voc = PascalVoc("2007", "trainval", str(tmp_path))
    db = voc.roidb()
    assert [r["uri"] for r in db] == ["000001", "000002"]
    np.testing.assert_allclose(db[0]["roi"], [[12, 10, 20, 100, 150, 0]])          # label 12 = dog, 0-based pixels, unknown class skipped
    from zoo.feature.image.imagePreprocessing import ImageResize
    from zoo.feature.image.roi import ImageRoiNormalize

    iset = voc.to_image_set()
    feats = [ImageResize(300, 300).apply(ImageRoiNormalize().apply(f)) for f in iset.features]
    x, t = roi_batch(feats)
    assert x.shape == (2, 3, 300, 300) and t.shape == (2, 7) and t[1, 0] == 1 and float(t[:, 2:6].max()) <= 1.0
