"""SSD object detection: prior boxes, box coding, the VGG-16 SSD network, the MultiBox training loss and the detection
output stage (decode + per-class NMS), as batched tensor programs.

Reference: zoo/.../models/image/objectdetection/ssd/SSD.scala:79-176 (the 300 / 512 configurations: prior sizes, aspect
ratios, steps, variances), ssd/SSDGraph.scala:58-215 (VGG base + extra layers + loc/conf heads + PriorBox),
common/loss/MultiBoxLoss.scala:30-37,112-260,332-397 (bipartite + per-prediction matching, hard negative mining 3:1,
SmoothL1 + softmax losses normalised by the number of matches), common/BboxUtil.scala (jaccard overlap, encode / decode
with variances, NMS), and BigDL's DetectionOutputSSD (conf 0.01, NMS 0.45, nmsTopk 400, keepTopK 200).

The reference walks Scala arrays box by box; here matching, mining and decoding are whole-batch tensor ops (one IoU matrix
per image, `topk` for mining), so the loss runs on the GPU next to the network.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch
import torch.nn as nn
import torch.nn.functional as F

VARIANCES = (0.1, 0.1, 0.2, 0.2)


# ------------------------------------------------------------------------------------------------ box utilities
def box_area(b):
    return (b[..., 2] - b[..., 0]).clamp(min=0) * (b[..., 3] - b[..., 1]).clamp(min=0)


def jaccard(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """IoU matrix of corner-form boxes a [A, 4] and b [B, 4] (BboxUtil.jaccardOverlap)."""
    lt = torch.max(a[:, None, :2], b[None, :, :2])
    rb = torch.min(a[:, None, 2:], b[None, :, 2:])
    inter = (rb - lt).clamp(min=0).prod(-1)
    union = box_area(a)[:, None] + box_area(b)[None, :] - inter
    return inter / union.clamp(min=1e-12)


def corner_to_center(b):
    return torch.cat([(b[..., :2] + b[..., 2:]) / 2, b[..., 2:] - b[..., :2]], -1)


def center_to_corner(b):
    return torch.cat([b[..., :2] - b[..., 2:] / 2, b[..., :2] + b[..., 2:] / 2], -1)


def encode_boxes(gt: torch.Tensor, priors: torch.Tensor, variances=VARIANCES) -> torch.Tensor:
    """Regression targets of corner-form gt boxes against corner-form priors (BboxUtil.encodeBBox, centre-size coding)."""
    g, p = corner_to_center(gt), corner_to_center(priors)
    v = gt.new_tensor(variances)
    d_xy = (g[..., :2] - p[..., :2]) / p[..., 2:] / v[:2]
    d_wh = torch.log((g[..., 2:] / p[..., 2:]).clamp(min=1e-12)) / v[2:]
    return torch.cat([d_xy, d_wh], -1)


def decode_boxes(loc: torch.Tensor, priors: torch.Tensor, variances=VARIANCES, clip=False) -> torch.Tensor:
    """Inverse of `encode_boxes` (BboxUtil.decodeBoxes)."""
    p = corner_to_center(priors)
    v = loc.new_tensor(variances)
    xy = loc[..., :2] * v[:2] * p[..., 2:] + p[..., :2]
    wh = torch.exp(loc[..., 2:] * v[2:]) * p[..., 2:]
    out = center_to_corner(torch.cat([xy, wh], -1))
    return out.clamp(0, 1) if clip else out


def nms(boxes: torch.Tensor, scores: torch.Tensor, iou_threshold=0.45, top_k=400) -> torch.Tensor:
    """Greedy non-maximum suppression; returns kept indices by decreasing score (BboxUtil.nms).  One IoU matrix for the
    `top_k` candidates, then a sequential sweep over a boolean mask (no per-box Python IoU)."""
    if boxes.numel() == 0:
        return torch.empty(0, dtype=torch.long, device=boxes.device)
    order = scores.argsort(descending=True)[:top_k]
    iou = jaccard(boxes[order], boxes[order])
    n = order.numel()
    suppressed = torch.zeros(n, dtype=torch.bool, device=boxes.device)
    over = (iou > iou_threshold).cpu()
    sup = suppressed.cpu()
    keep = []
    for i in range(n):
        if sup[i]:
            continue
        keep.append(i)
        sup |= over[i]
    return order[torch.tensor(keep, dtype=torch.long, device=boxes.device)]


# ------------------------------------------------------------------------------------------------ prior boxes
@dataclass
class PriorBoxSpec:
    """One feature map's prior boxes (SSDGraph.ComponetParam): sizes in input pixels."""
    feature_size: int
    min_size: float
    max_size: float
    aspect_ratios: tuple
    step: float
    flip: bool = True
    clip: bool = False

    @property
    def per_cell(self):
        return 2 + len(self.aspect_ratios) * (2 if self.flip else 1)


def prior_boxes(specs, resolution: int, offset=0.5) -> torch.Tensor:
    """Corner-form priors [P, 4] normalised to [0, 1], in the layout the heads produce (cell-major, then box)."""
    out = []
    for s in specs:
        f = s.feature_size
        ys, xs = torch.meshgrid(torch.arange(f, dtype=torch.float32), torch.arange(f, dtype=torch.float32), indexing="ij")
        cx = ((xs + offset) * s.step / resolution).reshape(-1, 1)
        cy = ((ys + offset) * s.step / resolution).reshape(-1, 1)
        whs = [(s.min_size, s.min_size), (math.sqrt(s.min_size * s.max_size),) * 2]
        for ar in s.aspect_ratios:
            r = math.sqrt(ar)
            whs.append((s.min_size * r, s.min_size / r))
            if s.flip:
                whs.append((s.min_size / r, s.min_size * r))
        wh = torch.tensor(whs, dtype=torch.float32) / resolution                      # [K, 2]
        k = wh.shape[0]
        c = torch.cat([cx, cy], 1)[:, None, :].expand(-1, k, -1)                       # [F*F, K, 2]
        b = torch.cat([c - wh[None] / 2, c + wh[None] / 2], -1).reshape(-1, 4)
        out.append(b.clamp(0, 1) if s.clip else b)
    return torch.cat(out, 0)


def ssd_specs(resolution=300, dataset="pascal", sizes=None):
    """The reference's two configurations (SSD.scala:104-176)."""
    if resolution not in (300, 512):
        raise ValueError("only support 300*300 or 512*512 as input")
    if sizes is None:
        table = {("pascal", 300): [30, 60, 111, 162, 213, 264, 315],
                 ("pascal", 512): [35.84, 76.8, 153.6, 230.4, 307.2, 384.0, 460.8, 537.6],
                 ("coco", 300): [21, 45, 99, 153, 207, 261, 315],
                 ("coco", 512): [20.48, 51.2, 133.12, 215.04, 296.96, 378.88, 460.8, 542.72]}
        if (dataset, resolution) not in table:
            raise NotImplementedError(f"{dataset} is not supported, please provide your own boundary sizes to parameter sizes")
        sizes = table[(dataset, resolution)]
    want = 7 if resolution == 300 else 8
    if len(sizes) != want:
        raise ValueError(f"the min and max division boundary length should be {want}")
    if resolution == 300:
        fmaps, steps, ars = [38, 19, 10, 5, 3, 1], [8, 16, 32, 64, 100, 300], [(2,), (2, 3), (2, 3), (2, 3), (2,), (2,)]
    else:
        fmaps, steps = [64, 32, 16, 8, 4, 2, 1], [8, 16, 32, 64, 128, 256, 512]
        ars = [(2,), (2, 3), (2, 3), (2, 3), (2, 3), (2,), (2,)]
    return [PriorBoxSpec(f, sizes[i], sizes[i + 1], ars[i], steps[i]) for i, f in enumerate(fmaps)]


# ------------------------------------------------------------------------------------------------ network
class L2NormScale(nn.Module):
    """Channel-wise L2 normalisation with a learned scale (conv4_3_norm, initial scale 20)."""

    def __init__(self, channels, scale=20.0):
        super().__init__()
        self.weight = nn.Parameter(torch.full((channels,), float(scale)))

    def forward(self, x):
        return F.normalize(x, p=2, dim=1, eps=1e-10) * self.weight.view(1, -1, 1, 1)


def _vgg16_base():
    cfg = [64, 64, "M", 128, 128, "M", 256, 256, 256, "C", 512, 512, 512]
    layers, c = [], 3
    for v in cfg:
        if v == "M":
            layers.append(nn.MaxPool2d(2, 2))
        elif v == "C":
            layers.append(nn.MaxPool2d(2, 2, ceil_mode=True))
        else:
            layers += [nn.Conv2d(c, v, 3, padding=1), nn.ReLU(inplace=True)]
            c = v
    conv4_3 = nn.Sequential(*layers)
    rest = nn.Sequential(nn.MaxPool2d(2, 2), nn.Conv2d(512, 512, 3, padding=1), nn.ReLU(inplace=True),
                         nn.Conv2d(512, 512, 3, padding=1), nn.ReLU(inplace=True), nn.Conv2d(512, 512, 3, padding=1), nn.ReLU(inplace=True),
                         nn.MaxPool2d(3, 1, padding=1), nn.Conv2d(512, 1024, 3, padding=6, dilation=6), nn.ReLU(inplace=True),   # fc6 (atrous)
                         nn.Conv2d(1024, 1024, 1), nn.ReLU(inplace=True))                                                   # fc7
    return conv4_3, rest


class SSDVGG(nn.Module):
    """SSD with a VGG-16 base.  `forward(x)` returns (loc [B, P, 4], conf [B, P, C]) in training mode and, in eval mode,
    detections [B, keep_top_k, 6] = (label, score, xmin, ymin, xmax, ymax) normalised to [0, 1] (label -1 = empty row)."""

    def __init__(self, class_num, resolution=300, dataset="pascal", sizes=None, nms_thresh=0.45, nms_topk=400, keep_topk=200,
                 conf_thresh=0.01, bg_label=0, base_channels_scale=1.0):
        super().__init__()
        self.class_num, self.resolution = int(class_num), int(resolution)
        self.specs = ssd_specs(resolution, dataset, sizes)
        self.conv4_3, self.fc7 = _vgg16_base()
        self.norm4_3 = L2NormScale(512, 20.0)
        ex = lambda cin, mid, cout, stride, pad, k=3: nn.Sequential(nn.Conv2d(cin, mid, 1), nn.ReLU(inplace=True),
                                                                  nn.Conv2d(mid, cout, k, stride=stride, padding=pad), nn.ReLU(inplace=True))
        if resolution == 300:
            self.extras = nn.ModuleList([ex(1024, 256, 512, 2, 1), ex(512, 128, 256, 2, 1), ex(256, 128, 256, 1, 0), ex(256, 128, 256, 1, 0)])
            chans = [512, 1024, 512, 256, 256, 256]
        else:
            self.extras = nn.ModuleList([ex(1024, 256, 512, 2, 1), ex(512, 128, 256, 2, 1), ex(256, 128, 256, 2, 1), ex(256, 128, 256, 2, 1),
                                         ex(256, 128, 256, 1, 1, k=4)])
            chans = [512, 1024, 512, 256, 256, 256, 256]
        self.loc = nn.ModuleList([nn.Conv2d(c, s.per_cell * 4, 3, padding=1) for c, s in zip(chans, self.specs)])
        self.conf = nn.ModuleList([nn.Conv2d(c, s.per_cell * self.class_num, 3, padding=1) for c, s in zip(chans, self.specs)])
        self.register_buffer("priors", prior_boxes(self.specs, resolution), persistent=False)
        self.detect = DetectionOutput(self.class_num, nms_thresh, nms_topk, keep_topk, conf_thresh, bg_label)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.xavier_uniform_(m.weight)
                nn.init.zeros_(m.bias)

    def features(self, x):
        f = self.conv4_3(x)
        maps = [self.norm4_3(f)]
        f = self.fc7(f)
        maps.append(f)
        for e in self.extras:
            f = e(f)
            maps.append(f)
        return maps

    def heads(self, x):
        locs, confs = [], []
        for f, l, c in zip(self.features(x), self.loc, self.conf):
            locs.append(l(f).permute(0, 2, 3, 1).reshape(x.shape[0], -1, 4))
            confs.append(c(f).permute(0, 2, 3, 1).reshape(x.shape[0], -1, self.class_num))
        return torch.cat(locs, 1), torch.cat(confs, 1)

    def forward(self, x):
        loc, conf = self.heads(x)
        if self.training:
            return loc, conf
        return self.detect(loc, conf, self.priors)


def SSD(class_num, resolution=300, dataset="pascal", sizes=None, **kw):
    """`SSDVGG(classNum, resolution, dataset, sizes, postProcessParam)` of the reference (SSD.scala:79-93)."""
    return SSDVGG(class_num, resolution, dataset, sizes, **kw)


# ------------------------------------------------------------------------------------------------ training loss
@dataclass
class MultiBoxLossParam:
    loc_weight: float = 1.0
    n_classes: int = 21
    share_location: bool = True
    overlap_threshold: float = 0.5
    bg_label_ind: int = 0
    use_difficult_gt: bool = True
    neg_pos_ratio: float = 3.0
    neg_overlap: float = 0.5


def match_priors(gt_boxes: torch.Tensor, priors: torch.Tensor, overlap_threshold=0.5):
    """Assign a ground-truth index (or -1) to every prior: each gt first claims its best prior (bipartite step), then every
    remaining prior takes its best gt if IoU > threshold (MultiBoxLoss.matchBbox).  Returns (match_idx [P], match_iou [P])."""
    P = priors.shape[0]
    if gt_boxes.numel() == 0:
        return torch.full((P,), -1, dtype=torch.long, device=priors.device), torch.zeros(P, device=priors.device)
    iou = jaccard(gt_boxes, priors)                                   # [G, P]
    best_iou, best_gt = iou.max(0)
    match = torch.where(best_iou > overlap_threshold, best_gt, torch.full_like(best_gt, -1))
    work = iou.clone()
    for _ in range(gt_boxes.shape[0]):                                # bipartite: highest remaining overlap first
        flat = work.argmax()
        g, p = int(flat // P), int(flat % P)
        if work[g, p] <= 1e-6:
            break
        match[p] = g
        best_iou[p] = iou[g, p]
        work[g, :] = -1
        work[:, p] = -1
    return match, best_iou


class MultiBoxLoss(nn.Module):
    """loss = (conf_loss + loc_weight * loc_loss) / num_matches.

    input : (loc [B, P, 4], conf [B, P, C]) and the priors [P, 4];
    target: a list of B tensors [G_i, 5+] = (label, xmin, ymin, xmax, ymax[, difficult]) normalised to [0, 1], or one tensor
            [N, 6+] = (image_index, label, xmin, ymin, xmax, ymax[, difficult]) like the reference's SSDMiniBatch target."""

    def __init__(self, param: MultiBoxLossParam | None = None, priors: torch.Tensor | None = None, **kw):
        super().__init__()
        self.param = param or MultiBoxLossParam(**kw)
        if priors is not None:
            self.register_buffer("priors", priors, persistent=False)
        else:
            self.priors = None

    @staticmethod
    def _split_target(target, batch):
        if isinstance(target, (list, tuple)):
            return list(target)
        out = []
        for b in range(batch):
            out.append(target[target[:, 0] == b][:, 1:])
        return out

    def forward(self, pred, target, priors=None):
        loc, conf = pred
        priors = priors if priors is not None else self.priors
        p = self.param
        B, P, C = conf.shape
        targets = self._split_target(target, B)
        loc_t = torch.zeros_like(loc)
        conf_t = torch.full((B, P), p.bg_label_ind, dtype=torch.long, device=conf.device)
        pos = torch.zeros((B, P), dtype=torch.bool, device=conf.device)
        neg_ok = torch.ones((B, P), dtype=torch.bool, device=conf.device)
        for b, t in enumerate(targets):
            t = t.to(loc.device)
            if t.numel() and not p.use_difficult_gt and t.shape[1] > 5:
                t = t[t[:, 5] == 0]
            if t.numel() == 0:
                continue
            match, miou = match_priors(t[:, 1:5].float(), priors, p.overlap_threshold)
            m = match >= 0
            pos[b] = m
            loc_t[b, m] = encode_boxes(t[match[m], 1:5].float(), priors[m]).to(loc.dtype)
            conf_t[b, m] = t[match[m], 0].long()
            neg_ok[b] = (~m) & (miou < p.neg_overlap)          # isEligibleMining
        num_pos = int(pos.sum())
        if num_pos == 0:
            return conf.sum() * 0.0
        # hard negative mining: per image, the neg_pos_ratio * num_pos eligible negatives with the highest background loss
        with torch.no_grad():
            logp = F.log_softmax(conf.float(), -1)
            bg_loss = -logp[..., p.bg_label_ind]
            bg_loss = bg_loss.masked_fill(~neg_ok, -1.0)
            n_neg = (pos.sum(1).float() * p.neg_pos_ratio).long().clamp(max=P - 1)
            order = bg_loss.argsort(1, descending=True)
            rank = order.argsort(1)
            neg = (rank < n_neg[:, None]) & neg_ok
        sel = pos | neg
        conf_loss = F.cross_entropy(conf[sel].float(), conf_t[sel], reduction="sum")
        loc_loss = F.smooth_l1_loss(loc[pos].float(), loc_t[pos].float(), reduction="sum")
        return (conf_loss + p.loc_weight * loc_loss) / num_pos


# ------------------------------------------------------------------------------------------------ detection output
class DetectionOutput(nn.Module):
    """Decode + softmax + per-class NMS + keep-top-k (BigDL DetectionOutputSSD as configured by SSD.scala:95-105)."""

    def __init__(self, n_classes, nms_thresh=0.45, nms_topk=400, keep_topk=200, conf_thresh=0.01, bg_label=0, variances=VARIANCES):
        super().__init__()
        self.n_classes, self.nms_thresh, self.nms_topk = n_classes, nms_thresh, nms_topk
        self.keep_topk, self.conf_thresh, self.bg_label, self.variances = keep_topk, conf_thresh, bg_label, variances

    @torch.no_grad()
    def forward(self, loc, conf, priors):
        B = loc.shape[0]
        scores = F.softmax(conf.float(), -1)
        boxes = decode_boxes(loc.float(), priors[None].float(), self.variances).clamp(0, 1)
        out = loc.new_full((B, self.keep_topk, 6), -1.0, dtype=torch.float32)
        for b in range(B):
            dets = []
            for c in range(self.n_classes):
                if c == self.bg_label:
                    continue
                s = scores[b, :, c]
                m = s > self.conf_thresh
                if not bool(m.any()):
                    continue
                bb, ss = boxes[b][m], s[m]
                keep = nms(bb, ss, self.nms_thresh, self.nms_topk)
                dets.append(torch.cat([torch.full((keep.numel(), 1), float(c), device=loc.device), ss[keep, None], bb[keep]], 1))
            if dets:
                d = torch.cat(dets, 0)
                d = d[d[:, 1].argsort(descending=True)[: self.keep_topk]]
                out[b, : d.shape[0]] = d
        return out
