"""Oracle: 3D detection postprocess. Test infrastructure only.

Torch-CPU restatements of the vendored pcdet functions the reference's Triton backend runs
(examples/pointpillar_kitti/1/model.py:163), each citing its source; rotated IoU/NMS come from
oracle/c/rotated_iou.c (pinned bit-for-bit against the reference's own compiled iou3d_cpu.cpp).
Tie rule fixed by the build wherever the reference leaves it to torch.topk/sort: score descending,
lower index first."""
import ctypes

import numpy as np
import torch

from .voxelize import clib


# --- rotated IoU / NMS (OPCD/pcdet/ops/iou3d_nms) -----------------------------------------------------
def boxes_iou_bev(a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    out = np.zeros((a.shape[0], b.shape[0]), np.float32)
    P = ctypes.c_void_p
    clib().oracle_boxes_iou_bev(P(a.ctypes.data), a.shape[0], P(b.ctypes.data), b.shape[0], P(out.ctypes.data))
    return out


def stable_topk(scores, k):
    """torch.topk with the build's tie rule (descending score, lower index first)."""
    s = torch.as_tensor(scores)
    order = torch.sort(s, descending=True, stable=True)[1][:k]
    return s[order], order


def nms_gpu(boxes, scores, thresh, pre_maxsize=None):
    """iou3d_nms_utils.nms_gpu (iou3d_nms_utils.py:84-99): sort descending, optional pre_maxsize cut, greedy sweep
    (iou3d_nms.cpp:116-132). Returns indices into `boxes` in kept order."""
    boxes = torch.as_tensor(boxes, dtype=torch.float32)
    scores = torch.as_tensor(scores, dtype=torch.float32)
    order = torch.sort(scores, descending=True, stable=True)[1]
    if pre_maxsize is not None:
        order = order[:pre_maxsize]
    sb = np.ascontiguousarray(boxes[order][:, :7].numpy(), np.float32)
    keep = np.zeros((max(1, sb.shape[0]),), np.int32)
    c_float = ctypes.c_float
    lib = clib()
    lib.oracle_rotated_nms_sorted.argtypes = [ctypes.c_void_p, ctypes.c_int, c_float, ctypes.c_void_p]
    nk = lib.oracle_rotated_nms_sorted(ctypes.c_void_p(sb.ctypes.data), sb.shape[0], float(thresh), ctypes.c_void_p(keep.ctypes.data))
    return order[torch.from_numpy(keep[:nk].astype(np.int64))]


def fma32(a, b, c):
    """float32 fma(a, b, c), exactly: the product is exact in float64, the sum is rounded to odd there (TwoSum gives the
    sign of what float64 dropped), and 53 bits >= 24 + 2 make the final rounding to float32 the correctly rounded one."""
    p = np.asarray(a, np.float32).astype(np.float64) * np.asarray(b, np.float32).astype(np.float64)
    c = np.asarray(c, np.float32).astype(np.float64)
    p, c = np.broadcast_arrays(p, c)
    s = p + c
    bb = s - p
    err = (p - (s - bb)) + (c - bb)
    odd = (s.view(np.int64) & 1).astype(bool)
    nxt = np.nextafter(s, np.where(err > 0, np.inf, -np.inf))
    with np.errstate(over="ignore"):
        return np.where((err != 0) & ~odd, nxt, s).astype(np.float32)


def iou_normal(b):
    """iou_normal (iou3d_nms_kernel.cu:314-325): axis-aligned BEV overlap of (x, y, dx, dy), heading ignored; float32.
    Element [i, j]: a = box i (the mask row), b = box j. This function exists only as CUDA in the reference; its nvcc
    build (12.9, sm_100a, SASS read back) contracts `Sa + Sb` into fma(b.dx, b.dy, Sa) with Sa = a.dx * a.dy rounded
    first and fuses nothing else that changes a value (x -+ dx / 2 as an fma by 0.5 is exact), so that is what is
    restated here."""
    b = np.ascontiguousarray(b, np.float32)
    f = np.float32
    x1, y1 = b[:, 0] - b[:, 3] / f(2), b[:, 1] - b[:, 4] / f(2)
    x2, y2 = b[:, 0] + b[:, 3] / f(2), b[:, 1] + b[:, 4] / f(2)
    w = np.maximum(np.minimum(x2[:, None], x2[None]) - np.maximum(x1[:, None], x1[None]), f(0))
    h = np.maximum(np.minimum(y2[:, None], y2[None]) - np.maximum(y1[:, None], y1[None]), f(0))
    inter = w * h
    area = b[:, 3] * b[:, 4]
    both = fma32(b[None, :, 3], b[None, :, 4], area[:, None])
    return inter / np.maximum(both - inter, f(1e-8))


def nms_normal_gpu(boxes, scores, thresh):
    """iou3d_nms_utils.nms_normal_gpu (iou3d_nms_utils.py:102-116): sort descending, greedy sweep with iou_normal."""
    boxes = torch.as_tensor(boxes, dtype=torch.float32)
    scores = torch.as_tensor(scores, dtype=torch.float32)
    order = torch.sort(scores, descending=True, stable=True)[1]
    over = iou_normal(boxes[order][:, :7].numpy()) > np.float32(thresh)
    n = over.shape[0]
    removed = np.zeros(n, bool)
    keep = []
    for i in range(n):
        if not removed[i]:
            keep.append(i)
            removed[i + 1:] |= over[i, i + 1:]
    return order[torch.tensor(keep, dtype=torch.int64)]


# --- reference: OPCD/pcdet/models/model_utils/model_nms_utils.py:6-25 ---------------------------------
def class_agnostic_nms(box_scores, box_preds, nms_thresh, pre_max, post_max, score_thresh=None, nms_type="nms_gpu"):
    src_box_scores = box_scores
    if score_thresh is not None:
        scores_mask = (box_scores >= score_thresh)
        box_scores = box_scores[scores_mask]
        box_preds = box_preds[scores_mask]
    selected = torch.zeros((0,), dtype=torch.int64)
    if box_scores.shape[0] > 0:
        box_scores_nms, indices = stable_topk(box_scores, min(pre_max, box_scores.shape[0]))
        boxes_for_nms = box_preds[indices]
        keep_idx = (nms_gpu if nms_type == "nms_gpu" else nms_normal_gpu)(boxes_for_nms[:, 0:7], box_scores_nms, nms_thresh)
        selected = indices[keep_idx[:post_max]]
    if score_thresh is not None:
        original_idxs = scores_mask.nonzero().view(-1)
        selected = original_idxs[selected]
    return selected, src_box_scores[selected]


# --- reference: OPCD/pcdet/models/model_utils/model_nms_utils.py:28-66 ---------------------------------
def multi_classes_nms(cls_scores, box_preds, nms_thresh, pre_max, post_max, score_thresh=None, nms_type="nms_gpu"):
    """Per class: score cut, top pre_max, NMS, first post_max; the classes' rows concatenated in class order."""
    cls_scores, box_preds = torch.as_tensor(cls_scores), torch.as_tensor(box_preds)
    pred_scores, pred_labels, pred_boxes = [], [], []
    for k in range(cls_scores.shape[1]):
        if score_thresh is not None:
            scores_mask = (cls_scores[:, k] >= score_thresh)
            box_scores = cls_scores[scores_mask, k]
            cur_box_preds = box_preds[scores_mask]
        else:
            box_scores = cls_scores[:, k]
            cur_box_preds = box_preds
        selected = torch.zeros((0,), dtype=torch.int64)
        if box_scores.shape[0] > 0:
            box_scores_nms, indices = stable_topk(box_scores, min(pre_max, box_scores.shape[0]))
            boxes_for_nms = cur_box_preds[indices]
            keep_idx = (nms_gpu if nms_type == "nms_gpu" else nms_normal_gpu)(boxes_for_nms[:, 0:7], box_scores_nms, nms_thresh)
            selected = indices[keep_idx[:post_max]]
        pred_scores.append(box_scores[selected])
        pred_labels.append(box_scores.new_ones(len(selected)).long() * k)
        pred_boxes.append(cur_box_preds[selected])
    return torch.cat(pred_scores, dim=0), torch.cat(pred_labels, dim=0), torch.cat(pred_boxes, dim=0)


# --- reference: OPCD/pcdet/models/dense_heads/target_assigner/anchor_generator.py:17-60 (CPU tensors) --
def generate_anchors(grid_size_xy, anchor_range, anchor_sizes, anchor_rotations, anchor_bottoms):
    """One entry per class; returns the concatenated [1, ny, nx, n_class, n_rot, 7] tensor of
    anchor_head_template.py:243 flattened to (-1, 7): order (y, x, class, rotation)."""
    anchor_range = np.asarray(anchor_range, np.float32)
    # pcdet passes numpy int64 feature-map sizes (anchor_head_template.py:30), so float32 / int64 promotes the strides
    # to float64. torch.arange then evaluates start + step*i in double in its scalar path and
    # float(start + step*i0) + lane*step inside SIMD chunks: the last ulp of an anchor coordinate depends on the CPU's
    # vector width. Tests therefore compare anchor x/y with a 1e-6 relative tolerance, everything else exactly.
    grid_size_xy = np.asarray(grid_size_xy, np.int64)
    all_anchors = []
    for anchor_size, anchor_rotation, anchor_height in zip(anchor_sizes, anchor_rotations, anchor_bottoms):
        x_stride = (anchor_range[3] - anchor_range[0]) / (grid_size_xy[0] - 1)
        y_stride = (anchor_range[4] - anchor_range[1]) / (grid_size_xy[1] - 1)
        x_shifts = torch.arange(anchor_range[0], anchor_range[3] + 1e-5, step=x_stride, dtype=torch.float32)
        y_shifts = torch.arange(anchor_range[1], anchor_range[4] + 1e-5, step=y_stride, dtype=torch.float32)
        z_shifts = x_shifts.new_tensor([anchor_height])
        num_anchor_size, num_anchor_rotation = 1, len(anchor_rotation)
        anchor_rotation = x_shifts.new_tensor(anchor_rotation)
        anchor_size_t = x_shifts.new_tensor([anchor_size])
        xs, ys, zs = torch.meshgrid([x_shifts, y_shifts, z_shifts], indexing="ij")
        anchors = torch.stack((xs, ys, zs), dim=-1)
        anchors = anchors[:, :, :, None, :].repeat(1, 1, 1, anchor_size_t.shape[0], 1)
        anchor_size_t = anchor_size_t.view(1, 1, 1, -1, 3).repeat([*anchors.shape[0:3], 1, 1])
        anchors = torch.cat((anchors, anchor_size_t), dim=-1)
        anchors = anchors[:, :, :, :, None, :].repeat(1, 1, 1, 1, num_anchor_rotation, 1)
        anchor_rotation = anchor_rotation.view(1, 1, 1, 1, -1, 1).repeat([*anchors.shape[0:3], num_anchor_size, 1, 1])
        anchors = torch.cat((anchors, anchor_rotation), dim=-1)
        anchors = anchors.permute(2, 1, 0, 3, 4, 5).contiguous()
        anchors[..., 2] += anchors[..., 5] / 2
        all_anchors.append(anchors)
    return torch.cat(all_anchors, dim=-3).view(-1, 7)


# --- reference: OPCD/pcdet/utils/box_coder_utils.py:45-77 (ResidualCoder.decode_torch) ----------------
def decode_torch(box_encodings, anchors):
    xa, ya, za, dxa, dya, dza, ra = torch.split(anchors, 1, dim=-1)
    xt, yt, zt, dxt, dyt, dzt, rt = torch.split(box_encodings, 1, dim=-1)
    diagonal = torch.sqrt(dxa ** 2 + dya ** 2)
    xg = xt * diagonal + xa
    yg = yt * diagonal + ya
    zg = zt * dza + za
    dxg = torch.exp(dxt) * dxa
    dyg = torch.exp(dyt) * dya
    dzg = torch.exp(dzt) * dza
    rg = rt + ra
    return torch.cat([xg, yg, zg, dxg, dyg, dzg, rg], dim=-1)


# --- reference: OPCD/pcdet/utils/common_utils.py:21-24 ------------------------------------------------
def limit_period(val, offset=0.5, period=np.pi):
    return val - torch.floor(val / period + offset) * period


# --- reference: OPCD/pcdet/models/dense_heads/anchor_head_template.py:225-272 -------------------------
def generate_predicted_boxes(cls_preds, box_preds, dir_cls_preds, anchors, dir_offset=0.78539, dir_limit_offset=0.0, num_dir_bins=2):
    num_anchors = anchors.shape[0]
    batch_cls_preds = torch.as_tensor(cls_preds).view(1, num_anchors, -1).float()
    batch_box_preds = decode_torch(torch.as_tensor(box_preds).view(1, num_anchors, -1), anchors.view(1, -1, 7))
    if dir_cls_preds is not None:
        dir_cls = torch.as_tensor(dir_cls_preds).view(1, num_anchors, -1)
        dir_labels = torch.max(dir_cls, dim=-1)[1]
        period = (2 * np.pi / num_dir_bins)
        dir_rot = limit_period(batch_box_preds[..., 6] - dir_offset, dir_limit_offset, period)
        batch_box_preds[..., 6] = dir_rot + dir_offset + period * dir_labels.to(batch_box_preds.dtype)
    return batch_cls_preds, batch_box_preds


# --- reference: OPCD/pcdet/models/detectors/detector3d_template.py:178-284 (MULTI_CLASSES_NMS False) --
def post_processing(batch_cls_preds, batch_box_preds, score_thresh=0.1, nms_thresh=0.01, pre_max=4096, post_max=500, index=0, nms_type="nms_gpu"):
    box_preds = batch_box_preds[index]
    cls_preds = torch.sigmoid(batch_cls_preds[index])
    cls_preds, label_preds = torch.max(cls_preds, dim=-1)
    label_preds = label_preds + 1
    selected, selected_scores = class_agnostic_nms(cls_preds, box_preds, nms_thresh, pre_max, post_max, score_thresh, nms_type)
    return {"pred_boxes": box_preds[selected], "pred_scores": selected_scores, "pred_labels": label_preds[selected]}


def pp_anchor_post(cls_preds, box_preds, dir_preds, anchor_range, anchor_sizes, anchor_bottoms, anchor_rotations, **kw):
    """Whole d2-d4 chain for one frame, composed from the pieces above."""
    h, w = cls_preds.shape[0], cls_preds.shape[1]
    anchors = generate_anchors((w, h), anchor_range, anchor_sizes, anchor_rotations, anchor_bottoms)
    post = {k: kw.pop(k) for k in ("score_thresh", "nms_thresh", "pre_max", "post_max", "nms_type") if k in kw}
    c, b = generate_predicted_boxes(cls_preds, box_preds, dir_preds, anchors, **kw)
    return post_processing(c, b, **post)


def pp_anchor_post_batch(cls_preds, box_preds, dir_preds, *a, **kw):
    """A batch [B,H,W,*]: the per-sample loop of Detector3DTemplate.post_processing (detector3d_template.py:196-199) —
    every sample is the single-frame chain on its own slice."""
    return [pp_anchor_post(cls_preds[b], box_preds[b], None if dir_preds is None else dir_preds[b], *a, **dict(kw)) for b in range(cls_preds.shape[0])]


# --- reference: OPCD/pcdet/models/model_utils/centernet_utils.py:118-151 ------------------------------
def _gather_feat(feat, ind):
    dim = feat.size(2)
    ind = ind.unsqueeze(2).expand(ind.size(0), ind.size(1), dim)
    return feat.gather(1, ind)


def _transpose_and_gather_feat(feat, ind):
    feat = feat.permute(0, 2, 3, 1).contiguous()
    feat = feat.view(feat.size(0), -1, feat.size(3))
    return _gather_feat(feat, ind)


def _topk(scores, K=40):
    batch, num_class, height, width = scores.size()
    flat = scores.flatten(2, 3)
    K1 = min(K, flat.shape[-1])
    order = torch.sort(flat, dim=-1, descending=True, stable=True)[1][..., :K1]
    topk_scores, topk_inds = flat.gather(-1, order), order
    topk_inds = topk_inds % (height * width)
    topk_ys = (topk_inds // width).float()
    topk_xs = (topk_inds % width).int().float()
    merged = topk_scores.view(batch, -1)
    K2 = min(K, merged.shape[-1])
    topk_ind = torch.sort(merged, dim=-1, descending=True, stable=True)[1][..., :K2]
    topk_score = merged.gather(-1, topk_ind)
    topk_classes = (topk_ind // K1).int()
    topk_inds = _gather_feat(topk_inds.view(batch, -1, 1), topk_ind).view(batch, K2)
    topk_ys = _gather_feat(topk_ys.view(batch, -1, 1), topk_ind).view(batch, K2)
    topk_xs = _gather_feat(topk_xs.view(batch, -1, 1), topk_ind).view(batch, K2)
    return topk_score, topk_inds, topk_classes, topk_ys, topk_xs


# --- reference: OPCD/pcdet/models/model_utils/centernet_utils.py:154-216 ------------------------------
def decode_bbox_from_heatmap(heatmap, rot_cos, rot_sin, center, center_z, dim, point_cloud_range, voxel_size,
                             feature_map_stride, vel=None, K=100, score_thresh=None, post_center_limit_range=None):
    batch_size = heatmap.shape[0]
    scores, inds, class_ids, ys, xs = _topk(heatmap, K=K)
    K = scores.shape[1]
    center = _transpose_and_gather_feat(center, inds).view(batch_size, K, 2)
    rot_sin = _transpose_and_gather_feat(rot_sin, inds).view(batch_size, K, 1)
    rot_cos = _transpose_and_gather_feat(rot_cos, inds).view(batch_size, K, 1)
    center_z = _transpose_and_gather_feat(center_z, inds).view(batch_size, K, 1)
    dim = _transpose_and_gather_feat(dim, inds).view(batch_size, K, 3)
    angle = torch.atan2(rot_sin, rot_cos)
    xs = xs.view(batch_size, K, 1) + center[:, :, 0:1]
    ys = ys.view(batch_size, K, 1) + center[:, :, 1:2]
    xs = xs * feature_map_stride * voxel_size[0] + point_cloud_range[0]
    ys = ys * feature_map_stride * voxel_size[1] + point_cloud_range[1]
    box_part_list = [xs, ys, center_z, dim, angle]
    if vel is not None:
        box_part_list.append(_transpose_and_gather_feat(vel, inds).view(batch_size, K, 2))
    final_box_preds = torch.cat(box_part_list, dim=-1)
    final_scores = scores.view(batch_size, K)
    final_class_ids = class_ids.view(batch_size, K)
    mask = (final_box_preds[..., :3] >= post_center_limit_range[:3]).all(2)
    mask &= (final_box_preds[..., :3] <= post_center_limit_range[3:]).all(2)
    if score_thresh is not None:
        mask &= (final_scores > score_thresh)
    return [{"pred_boxes": final_box_preds[k, mask[k]], "pred_scores": final_scores[k, mask[k]], "pred_labels": final_class_ids[k, mask[k]]}
            for k in range(batch_size)]


# --- reference: OPCD/pcdet/models/dense_heads/center_head.py:253-304 ----------------------------------
def centerhead_generate_predicted_boxes(heads, class_id_mapping_each_head, point_cloud_range, voxel_size, feature_map_stride,
                                        post_center_limit_range, K=500, score_thresh=0.1, nms_thresh=0.2, pre_max=1000, post_max=83,
                                        nms_type="nms_gpu"):
    limit = torch.tensor(post_center_limit_range).float()
    boxes, scores, labels = [], [], []
    for idx, pd in enumerate(heads):
        t = {k: torch.as_tensor(v)[None] for k, v in pd.items()}
        batch_hm = t["hm"].sigmoid()
        batch_dim = t["dim"].exp()
        final = decode_bbox_from_heatmap(
            heatmap=batch_hm, rot_cos=t["rot"][:, 0].unsqueeze(dim=1), rot_sin=t["rot"][:, 1].unsqueeze(dim=1), center=t["center"],
            center_z=t["center_z"], dim=batch_dim, vel=t.get("vel"), point_cloud_range=point_cloud_range, voxel_size=voxel_size,
            feature_map_stride=feature_map_stride, K=K, score_thresh=score_thresh, post_center_limit_range=limit)[0]
        final["pred_labels"] = torch.as_tensor(class_id_mapping_each_head[idx])[final["pred_labels"].long()]
        selected, selected_scores = class_agnostic_nms(final["pred_scores"], final["pred_boxes"], nms_thresh, pre_max, post_max, None, nms_type)
        boxes.append(final["pred_boxes"][selected])
        scores.append(selected_scores)
        labels.append(final["pred_labels"][selected])
    return {"pred_boxes": torch.cat(boxes, 0), "pred_scores": torch.cat(scores, 0), "pred_labels": torch.cat(labels, 0) + 1}


def centerhead_generate_predicted_boxes_batch(heads, *a, **kw):
    """heads: per head a dict of [B,c,H,W] arrays; CenterHead.generate_predicted_boxes(batch_size, ...) keeps the samples
    apart (ret_dict[k], center_head.py:257-300), so sample b is the single-frame chain on slice b."""
    B = next(iter(heads[0].values())).shape[0]
    return [centerhead_generate_predicted_boxes([{k: v[b] for k, v in h.items()} for h in heads], *a, **dict(kw)) for b in range(B)]


# --- reference: clients/postprocess/detector_3d_postprocess.py:55-64, 98-133 --------------------------
NUSC_GROUPS = [(0, 0.4), (8, 0.1), (7, 0.15), (3, 0.3), (2, 0.4), (9, 0.1), (4, 0.4), (5, 0.4), (1, 0.4)]
# car, pedestrian, bicycle, bus, construction_vehicle, traffic_cone, trailer, barrier, truck — the reference
# computes motorcycle (6, 0.15) indices too but never appends them (:118-127).


def get_annotations_indices(types, thresh, label_preds, scores):
    indexs = [i for i in range(label_preds.shape[0]) if label_preds[i] == types]
    return [i for i in indexs if scores[i] >= thresh]


def remove_low_score_nu(predictions):
    labels, scores = predictions["labels"], predictions["scores"]
    idx = []
    for cls, thr in NUSC_GROUPS:
        idx += get_annotations_indices(cls, thr, labels, scores)
    return {k: v[idx] for k, v in predictions.items() if k != "metadata"}
