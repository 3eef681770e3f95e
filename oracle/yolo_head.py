"""Oracle: YOLO raw-head decode and the YOLOv4 two-tensor postprocess. Test infrastructure only."""
import numpy as np
import torch

from .yolo_post import nms_cpu


# --- reference: tools/yolo_layer.py:148-288 (yolo_forward_dynamic) ------------------------------------
def yolo_forward_dynamic(output, num_classes, anchors, num_anchors, scale_x_y=1):
    """output [B, A*(5+nc), H, W] float32; anchors already divided by the stride (YoloLayer.forward :313-321).
    Returns boxes [B, A*H*W, 1, 4], confs [B, A*H*W, nc]. Same torch ops in the same order as the reference."""
    output = torch.as_tensor(output)
    B, _, H, W = output.shape
    bxy_list, bwh_list, det_list, cls_list = [], [], [], []
    for i in range(num_anchors):
        begin, end = i * (5 + num_classes), (i + 1) * (5 + num_classes)
        bxy_list.append(output[:, begin:begin + 2])
        bwh_list.append(output[:, begin + 2:begin + 4])
        det_list.append(output[:, begin + 4:begin + 5])
        cls_list.append(output[:, begin + 5:end])
    bxy, bwh = torch.cat(bxy_list, dim=1), torch.cat(bwh_list, dim=1)
    det_confs = torch.cat(det_list, dim=1).view(B, num_anchors * H * W)
    cls_confs = torch.cat(cls_list, dim=1).view(B, num_anchors, num_classes, H * W)
    cls_confs = cls_confs.permute(0, 1, 3, 2).reshape(B, num_anchors * H * W, num_classes)
    bxy = torch.sigmoid(bxy) * scale_x_y - 0.5 * (scale_x_y - 1)
    bwh = torch.exp(bwh)
    det_confs = torch.sigmoid(det_confs)
    cls_confs = torch.sigmoid(cls_confs)
    grid_x = np.expand_dims(np.expand_dims(np.expand_dims(np.linspace(0, W - 1, W), axis=0).repeat(H, 0), axis=0), axis=0)
    grid_y = np.expand_dims(np.expand_dims(np.expand_dims(np.linspace(0, H - 1, H), axis=1).repeat(W, 1), axis=0), axis=0)
    bx_list, by_list, bw_list, bh_list = [], [], [], []
    for i in range(num_anchors):
        ii = i * 2
        bx_list.append(bxy[:, ii:ii + 1] + torch.tensor(grid_x, dtype=torch.float32))
        by_list.append(bxy[:, ii + 1:ii + 2] + torch.tensor(grid_y, dtype=torch.float32))
        bw_list.append(bwh[:, ii:ii + 1] * anchors[i * 2])
        bh_list.append(bwh[:, ii + 1:ii + 2] * anchors[i * 2 + 1])
    bx_bw = torch.cat((torch.cat(bx_list, dim=1), torch.cat(bw_list, dim=1)), dim=1)
    by_bh = torch.cat((torch.cat(by_list, dim=1), torch.cat(bh_list, dim=1)), dim=1)
    bx_bw /= W
    by_bh /= H
    n = num_anchors * H * W
    bx, by = bx_bw[:, :num_anchors].reshape(B, n, 1), by_bh[:, :num_anchors].reshape(B, n, 1)
    bw, bh = bx_bw[:, num_anchors:].reshape(B, n, 1), by_bh[:, num_anchors:].reshape(B, n, 1)
    bx1, by1 = bx - bw * 0.5, by - bh * 0.5
    bx2, by2 = bx1 + bw, by1 + bh
    boxes = torch.cat((bx1, by1, bx2, by2), dim=2).view(B, n, 1, 4)
    confs = cls_confs * det_confs.view(B, n, 1)
    return boxes, confs


# --- reference: tools/torch_utils.py:48-64 (get_region_boxes) -----------------------------------------
def get_region_boxes(boxes_and_confs):
    return torch.cat([b for b, _ in boxes_and_confs], dim=1), torch.cat([c for _, c in boxes_and_confs], dim=1)


# --- NOT in the reference: ultralytics YOLOv5 Detect decode, parity unpinned -----------------
def yolov5_detect(heads, anchors_per_head, strides, num_classes):
    """heads [B, A*(5+nc), H, W] -> [B, sum(A*H*W), 5+nc], rows ordered (anchor, y, x) per head."""
    z = []
    no = 5 + num_classes
    for x, anchors, s in zip(heads, anchors_per_head, strides):
        x = torch.as_tensor(x)
        B, _, H, W = x.shape
        na = len(anchors) // 2
        y = x.view(B, na, no, H, W).permute(0, 1, 3, 4, 2).contiguous().sigmoid()
        yv, xv = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
        grid = torch.stack((xv, yv), 2).view(1, 1, H, W, 2).float()
        ag = torch.tensor(anchors, dtype=torch.float32).view(1, na, 1, 1, 2)
        xy = (y[..., 0:2] * 2. - 0.5 + grid) * s
        wh = (y[..., 2:4] * 2) ** 2 * ag
        z.append(torch.cat((xy, wh, y[..., 4:]), -1).view(B, -1, no))
    return torch.cat(z, 1)


# --- reference: utils/postprocess.py:265-322 (extract_boxes_onnx; tools/utils.py:166-233 is the same) ----
def extract_boxes_onnx(conf_thresh, nms_thresh, output, stable=True):
    box_array, confs = np.asarray(output[0]), np.asarray(output[1])
    num_classes = confs.shape[2]
    box_array = box_array[:, :, 0]
    max_conf = np.max(confs, axis=2)
    max_id = np.argmax(confs, axis=2)
    bboxes_batch = []
    for i in range(box_array.shape[0]):
        argwhere = max_conf[i] > conf_thresh
        l_box_array, l_max_conf, l_max_id = box_array[i, argwhere, :], max_conf[i, argwhere], max_id[i, argwhere]
        bboxes = []
        for j in range(num_classes):
            cls_argwhere = l_max_id == j
            ll_box_array, ll_max_conf, ll_max_id = l_box_array[cls_argwhere, :], l_max_conf[cls_argwhere], l_max_id[cls_argwhere]
            keep = nms_cpu(ll_box_array, ll_max_conf, nms_thresh, stable=stable)
            if keep.size > 0:
                ll_box_array, ll_max_conf, ll_max_id = ll_box_array[keep, :], ll_max_conf[keep], ll_max_id[keep]
                for k in range(ll_box_array.shape[0]):
                    bboxes.append([ll_box_array[k, 0], ll_box_array[k, 1], ll_box_array[k, 2], ll_box_array[k, 3],
                                   ll_max_conf[k], ll_max_conf[k], ll_max_id[k]])
        bboxes_batch.append(bboxes)
    return bboxes_batch
