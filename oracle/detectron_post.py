"""Oracle: detectron2 dense-detector inference (RetinaNet / FCOS), SURVEY §8 row c7 / Appendix A6.
Test infrastructure only. detectron2 is not vendored by the reference and no version is pinned -> PARITY UNPINNED;
this restates its published per-level inference with torch ops and uses torchvision.ops.batched_nms as the NMS."""
import math

import numpy as np
import torch
import torchvision


def cell_anchors(sizes, ratios):
    """DefaultAnchorGenerator.generate_cell_anchors: for size, for ratio: w = sqrt(size^2/ratio), h = ratio*w."""
    out = []
    for size in sizes:
        area = size ** 2.0
        for r in ratios:
            w = math.sqrt(area / r)
            h = r * w
            out.append([-w / 2.0, -h / 2.0, w / 2.0, h / 2.0])
    return out


def retinanet_cell_anchors(levels=5):
    return [cell_anchors([32 * 2 ** l * s for s in (1, 2 ** (1 / 3), 2 ** (2 / 3))], (0.5, 1.0, 2.0)) for l in range(levels)]


def fcos_cell_anchors(strides):
    return [cell_anchors([s], (1.0,)) for s in strides]


def grid_anchors(h, w, stride, cell):
    ys, xs = torch.meshgrid(torch.arange(0, h * stride, stride, dtype=torch.float32), torch.arange(0, w * stride, stride, dtype=torch.float32), indexing="ij")
    shifts = torch.stack((xs.reshape(-1), ys.reshape(-1), xs.reshape(-1), ys.reshape(-1)), dim=1)
    return (shifts.view(-1, 1, 4) + torch.tensor(cell, dtype=torch.float32).view(1, -1, 4)).reshape(-1, 4)


def apply_deltas(deltas, boxes):
    widths, heights = boxes[:, 2] - boxes[:, 0], boxes[:, 3] - boxes[:, 1]
    ctr_x, ctr_y = boxes[:, 0] + 0.5 * widths, boxes[:, 1] + 0.5 * heights
    dx, dy, dw, dh = deltas[:, 0], deltas[:, 1], deltas[:, 2], deltas[:, 3]
    clamp = math.log(1000.0 / 16)
    dw, dh = torch.clamp(dw, max=clamp), torch.clamp(dh, max=clamp)
    pcx, pcy = dx * widths + ctr_x, dy * heights + ctr_y
    pw, ph = torch.exp(dw) * widths, torch.exp(dh) * heights
    return torch.stack((pcx - 0.5 * pw, pcy - 0.5 * ph, pcx + 0.5 * pw, pcy + 0.5 * ph), dim=1)


def apply_deltas_linear(deltas, boxes):
    widths, heights = boxes[:, 2] - boxes[:, 0], boxes[:, 3] - boxes[:, 1]
    ctr_x, ctr_y = (boxes[:, 0] + boxes[:, 2]) / 2, (boxes[:, 1] + boxes[:, 3]) / 2
    d = torch.relu(deltas)
    return torch.stack((ctr_x - d[:, 0] * widths, ctr_y - d[:, 1] * heights, ctr_x + d[:, 2] * widths, ctr_y + d[:, 3] * heights), dim=1)


def dense_inference(logits, deltas, strides, cells, image_hw, ctrness=None, score_thresh=0.05, topk=1000, nms_thresh=0.5, max_det=100):
    boxes_all, scores_all, cls_all = [], [], []
    for l, (lg, dl) in enumerate(zip(logits, deltas)):
        lg, dl = torch.as_tensor(lg), torch.as_tensor(dl)
        ak, h, w = lg.shape
        a = len(cells[l])
        k = ak // a
        s = lg.view(a, k, h, w).permute(2, 3, 0, 1).reshape(-1, k).sigmoid()            # (H*W*A, K)
        if ctrness is not None:
            c = torch.as_tensor(ctrness[l]).view(1, h, w).permute(1, 2, 0).reshape(-1, 1).sigmoid()
            s = torch.sqrt(s.view(h * w, a, k) * c.view(h * w, 1, 1)).reshape(-1, k)
        d = dl.view(a, 4, h, w).permute(2, 3, 0, 1).reshape(-1, 4)
        flat = s.reshape(-1)
        keep = torch.nonzero(flat > score_thresh).view(-1)
        sc = flat[keep]
        order = torch.sort(sc, descending=True, stable=True)[1][:min(topk, sc.numel())]
        sc, idx = sc[order], keep[order]
        anchor_idx, cls_idx = idx // k, idx % k
        anchors = grid_anchors(h, w, strides[l], cells[l])
        bx = (apply_deltas if ctrness is None else apply_deltas_linear)(d[anchor_idx], anchors[anchor_idx])
        boxes_all.append(bx)
        scores_all.append(sc)
        cls_all.append(cls_idx)
    boxes, scores, classes = torch.cat(boxes_all), torch.cat(scores_all), torch.cat(cls_all)
    boxes[:, 0::2] = boxes[:, 0::2].clamp(0, image_hw[1])
    boxes[:, 1::2] = boxes[:, 1::2].clamp(0, image_hw[0])
    # torchvision.ops.batched_nms picks its strategy by device (coordinate trick up to 1000 boxes on CPU, 5000 on CUDA).
    # The served model runs on CUDA, so the oracle spells the coordinate trick out (boxes.py: _batched_nms_coordinate_trick).
    if boxes.numel() == 0:
        keep = torch.zeros((0,), dtype=torch.int64)
    else:
        max_coordinate = boxes.max()
        offsets = classes.to(boxes) * (max_coordinate + torch.tensor(1).to(boxes))
        keep = torchvision.ops.nms(boxes + offsets[:, None], scores, nms_thresh)[:max_det]
    return boxes[keep].numpy(), scores[keep].numpy(), classes[keep].numpy()
