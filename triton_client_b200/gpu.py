"""Device-level Python wrappers: torch CUDA tensors in, torch CUDA tensors out, every op one call
through the C ABI of libtcb200.so. No op here has a CPU or eager-PyTorch fallback.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import c_float, c_int, c_void_p

STRETCH, LETTERBOX, PAD = 0, 1, 2
_MODES = {"stretch": STRETCH, "letterbox": LETTERBOX, "pad": PAD}


def _ptr(t):
    return c_void_p(t.data_ptr())


def _need_cuda(*tensors):
    for t in tensors:
        if not (isinstance(t, torch.Tensor) and t.is_cuda):
            raise TypeError("triton_client_b200.gpu ops take CUDA tensors only (there is no CPU path)")


def _nms_type(name):
    """NMS_CONFIG.NMS_TYPE -> TCB_NMS_GPU / TCB_NMS_NORMAL; the reference resolves the name with
    getattr(iou3d_nms_utils, NMS_TYPE) (model_nms_utils.py:15), so any other name is an AttributeError there too."""
    try:
        return {"nms_gpu": 0, "nms_normal_gpu": 1}[name]
    except KeyError:
        raise AttributeError(f"module 'iou3d_nms_utils' has no attribute {name!r}") from None


class Engine:
    """A tcb_ctx plus the torch plumbing around it. One per pre/post object (one frame in flight)."""

    def __init__(self, device=0, workspace_bytes=256 << 20):
        if not torch.cuda.is_available():
            raise RuntimeError("triton_client_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", int(device))
        torch.cuda.set_device(self.device)
        torch.zeros(1, device=self.device)  # make sure the primary context exists before the library uses it
        self.ctx = _lib.Context(int(device), int(workspace_bytes))
        self.lib = _lib.lib()

    @property
    def launches(self):
        return self.ctx.launches

    # ---- 2D preprocess ---------------------------------------------------------------------------
    def preprocess_images(self, src, out_h, out_w, mode="stretch", swap=False, div=255.0, sub=None,
                          pad_value=114, out=None):
        """src: uint8 CUDA tensor [n,h,w,3] (or [h,w,3]), any row/frame stride with contiguous pixels.
        Returns (f32 [n,3,out_h,out_w], inv_affine[4] = gain_x, gain_y, pad_left, pad_top)."""
        _need_cuda(src)
        if src.dtype != torch.uint8:
            raise TypeError("preprocess_images: src must be uint8")
        if src.dim() == 3:
            src = src.unsqueeze(0)
        n, h, w, c = src.shape
        if c != 3 or src.stride(3) != 1 or src.stride(2) != 3:
            raise ValueError("preprocess_images: need packed 3-channel pixels (HWC, 3 bytes per pixel)")
        if out is None:
            out = torch.empty((n, 3, out_h, out_w), dtype=torch.float32, device=src.device)
        elif tuple(out.shape) != (n, 3, out_h, out_w) or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError("preprocess_images: bad `out` tensor")
        sub_arr = (c_float * 3)(*sub) if sub is not None else None       # NULL = no subtraction
        aff = (c_float * 4)()
        self.ctx.call("tcb_preprocess_image", _ptr(src), n, h, w, src.stride(1), src.stride(0) if n > 1 else h * src.stride(1),
                      1 if swap else 0, _MODES[mode] if isinstance(mode, str) else int(mode), out_h, out_w,
                      float(div), sub_arr, int(pad_value), _ptr(out), aff, _lib.stream_ptr())
        return out, np.frombuffer(aff, dtype=np.float32)     # the array keeps `aff` alive

    # ---- YOLOv5 postprocess ----------------------------------------------------------------------
    # ---- compressed camera frames ---------------------------------------------------------------------
    def jpeg_info(self, data):
        """(height, width, channels) of a baseline JPEG stream (bytes-like)."""
        buf = np.frombuffer(data, np.uint8)
        h, w, c = c_int(), c_int(), c_int()
        rc = self.lib.tcb_jpeg_info(c_void_p(buf.ctypes.data), buf.size, ctypes.byref(h), ctypes.byref(w), ctypes.byref(c))
        if rc != 0:
            raise _lib.TcbError(rc, (self.lib.tcb_last_error(None) or b"").decode())
        return h.value, w.value, c.value

    def jpeg_decode(self, streams, bgr=False, out=None):
        """cv2.imdecode(..., IMREAD_ANYCOLOR) of a list of same-sized baseline JPEG streams (bytes-like), on the device:
        uint8 CUDA tensor [n, H, W, 3] in RGB order (the reference flips cv2's BGR to RGB right away; bgr=True keeps
        cv2's order) or [n, H, W] for grey streams. Bit-identical to OpenCV / libjpeg-turbo."""
        bufs = [np.frombuffer(s, np.uint8) for s in streams]
        n = len(bufs)
        if n == 0:
            return torch.empty((0, 0, 0, 3), dtype=torch.uint8, device=self.device)
        h, w, c = self.jpeg_info(bufs[0])
        total = sum(b.size for b in bufs)
        need = int(self.lib.tcb_jpeg_staging_bytes(total, n))
        # two pinned staging buffers taken in turn: the host parse of this call overlaps the kernels of the previous one,
        # and a buffer is rewritten only after the upload that read it (two calls ago) has finished
        if getattr(self, "_jpeg_slots", None) is None:
            self._jpeg_slots, self._jpeg_turn = [[None, None], [None, None]], 0
        slot = self._jpeg_slots[self._jpeg_turn]
        self._jpeg_turn ^= 1
        if slot[1] is not None:
            slot[1].synchronize()
        if slot[0] is None or slot[0].numel() < need:
            slot[0] = torch.empty(max(need, 1 << 20), dtype=torch.uint8).pin_memory()
        if out is None:
            out = torch.empty((n, h, w, 3) if c == 3 else (n, h, w), dtype=torch.uint8, device=self.device)
        ptrs = (c_void_p * n)(*[b.ctypes.data for b in bufs])
        lens = (ctypes.c_size_t * n)(*[b.size for b in bufs])
        self.ctx.call("tcb_jpeg_decode", ptrs, lens, n, c_void_p(slot[0].data_ptr()), slot[0].numel(), 1 if bgr else 0,
                      _ptr(out), h, w, c, _lib.stream_ptr())
        slot[1] = torch.cuda.Event()
        slot[1].record()
        return out

    MODES = {"stretch": 0, "letterbox": 1, "pad": 2}

    def content_rect(self, mode, h, w, out_h, out_w):
        """(left, top, width, height) of the resampled frame inside the (out_h, out_w) output of preprocess_images."""
        rect = (c_int * 4)()
        rc = self.lib.tcb_content_rect(self.MODES[mode], int(h), int(w), int(out_h), int(out_w), rect)
        if rc != 0:
            raise _lib.TcbError(rc, (self.lib.tcb_last_error(None) or b"").decode())
        return tuple(int(v) for v in rect)

    def copy_rect_d2h(self, src, dst_host, rect):
        """Asynchronous download (current stream) of rectangle `rect` of every [H,W] plane of CUDA tensor `src` [..., H, W]
        into the pinned host tensor `dst_host` of the same shape; the rest of `dst_host` is left as it is."""
        _need_cuda(src)
        if not (src.is_contiguous() and dst_host.is_contiguous() and dst_host.is_pinned() and tuple(src.shape) == tuple(dst_host.shape)
                and src.dtype == dst_host.dtype):
            raise TypeError("copy_rect_d2h: need a contiguous CUDA source and a pinned, contiguous host destination of the same shape and dtype")
        h, w = src.shape[-2], src.shape[-1]
        planes = src.numel() // (h * w) if h * w else 0
        left, top, rw, rh = rect
        self.ctx.call("tcb_copy_rect_d2h", _ptr(src), c_void_p(dst_host.data_ptr()), planes, h, w, src.element_size(), left, top, rw, rh,
                      _lib.stream_ptr())

    def yolo_nms_workspace(self, b, n):
        return int(self.lib.tcb_yolo_nms_workspace(int(b), int(n)))

    def yolo_nms(self, pred, conf_thres=0.6, iou_thres=0.45, classes=None, agnostic=False, max_det=300,
                 max_nms=30000, out=None, counts=None):
        """pred: f32 CUDA tensor [b,n,5+nc]. Returns (out f64 [b,max_det,6], counts i32 [b], xc i32 [b])."""
        _need_cuda(pred)
        if pred.dtype != torch.float32 or pred.dim() != 3 or not pred.is_contiguous():
            raise TypeError("yolo_nms: pred must be a contiguous float32 [b,n,5+nc] tensor")
        b, n, c = pred.shape
        if out is None:
            out = torch.zeros((b, max_det, 6), dtype=torch.float64, device=pred.device)
        if counts is None:
            counts = torch.zeros((2, b), dtype=torch.int32, device=pred.device)
        cls_arr, ncls = None, 0
        if classes is not None:
            ncls = len(classes)
            cls_arr = (c_int * max(ncls, 1))(*[int(v) for v in classes])
        cp = counts.data_ptr()                         # rows of the [2, b] int32 tensor, without building views for the call
        self.ctx.call("tcb_yolo_nms", _ptr(pred), b, n, c - 5, float(conf_thres), float(iou_thres), cls_arr, ncls,
                      1 if agnostic else 0, int(max_det), int(max_nms), _ptr(out), c_void_p(cp), c_void_p(cp + 4 * counts.stride(0)),
                      _lib.stream_ptr())
        return out, counts[0], counts[1]

    # ---- YOLO head decode, YOLOv4 post, axis-aligned helpers ------------------------------------------
    def yolo_head_decode_v4(self, heads, anchors_per_head, num_classes, scale_x_y=1.0):
        """heads: list of CUDA f32 [B, A*(5+nc), H, W]; anchors_per_head: list of flat [w0,h0,w1,h1,...] already divided by
        the stride (YoloLayer.forward). Returns (boxes [B,N,1,4], confs [B,N,nc]) with the heads concatenated."""
        b = heads[0].shape[0]
        rows = [len(a) // 2 * h.shape[2] * h.shape[3] for h, a in zip(heads, anchors_per_head)]
        total = sum(rows)
        boxes = torch.empty((b, total, 1, 4), dtype=torch.float32, device=heads[0].device)
        confs = torch.empty((b, total, num_classes), dtype=torch.float32, device=heads[0].device)
        off = 0
        for h, a, r in zip(heads, anchors_per_head, rows):
            _need_cuda(h)
            arr = (c_float * len(a))(*[float(v) for v in a])
            self.ctx.call("tcb_yolo_head_decode", _ptr(h.contiguous()), b, h.shape[2], h.shape[3], len(a) // 2, num_classes, arr, 0,
                          float(scale_x_y), 1.0, off, total, _ptr(boxes), _ptr(confs), None, _lib.stream_ptr())
            off += r
        return boxes, confs

    def yolo_head_decode_v5(self, heads, anchors_per_head, strides, num_classes):
        """YOLOv5 Detect: heads [B, A*(5+nc), H, W] -> pred [B, N, 5+nc] (the input of yolo_nms); anchors in pixels."""
        b = heads[0].shape[0]
        rows = [len(a) // 2 * h.shape[2] * h.shape[3] for h, a in zip(heads, anchors_per_head)]
        total = sum(rows)
        pred = torch.empty((b, total, 5 + num_classes), dtype=torch.float32, device=heads[0].device)
        off = 0
        for h, a, st, r in zip(heads, anchors_per_head, strides, rows):
            _need_cuda(h)
            arr = (c_float * len(a))(*[float(v) for v in a])
            self.ctx.call("tcb_yolo_head_decode", _ptr(h.contiguous()), b, h.shape[2], h.shape[3], len(a) // 2, num_classes, arr, 1,
                          1.0, float(st), off, total, None, None, _ptr(pred), _lib.stream_ptr())
            off += r
        return pred

    def nms_axis_aligned(self, boxes, scores, thresh=0.5, min_mode=False, post_max=1024):
        """nms_cpu on the GPU: returns (keep int32 [post_max], count int32 [2] = kept, overflow flag)."""
        _need_cuda(boxes, scores)
        b, s = boxes.contiguous().float(), scores.contiguous().float()
        keep = torch.empty((post_max,), dtype=torch.int32, device=b.device)
        cnt = torch.zeros((2,), dtype=torch.int32, device=b.device)
        self.ctx.call("tcb_nms_axis_aligned", _ptr(b), _ptr(s), b.shape[0], float(thresh), 1 if min_mode else 0, int(post_max),
                      _ptr(keep), _ptr(cnt), _lib.stream_ptr())
        return keep, cnt

    def yolo4_post(self, boxes, confs, conf_thresh=0.4, nms_thresh=0.6, post_max=1024):
        """extract_boxes_onnx: boxes [B,N,1,4] (or [B,N,4]), confs [B,N,nc] -> (rows [B,post_max,7], counts [2,B]:
        kept rows per image, and 1 where more than 4096 rows passed the confidence cut (result incomplete))."""
        _need_cuda(boxes, confs)
        b, n, nc = confs.shape
        bx = boxes.reshape(b, n, 4).contiguous().float()
        out = torch.zeros((b, post_max, 7), dtype=torch.float32, device=bx.device)
        cnt = torch.zeros((2, b), dtype=torch.int32, device=bx.device)
        self.ctx.call("tcb_yolo4_post", _ptr(bx), _ptr(confs.contiguous().float()), b, n, nc, float(conf_thresh), float(nms_thresh),
                      int(post_max), _ptr(out), _ptr(cnt), _lib.stream_ptr())
        return out, cnt

    def box_iou(self, box1, box2):
        _need_cuda(box1, box2)
        a, b = box1.contiguous().float(), box2.contiguous().float()
        out = torch.empty((a.shape[0], b.shape[0]), dtype=torch.float32, device=a.device)
        self.ctx.call("tcb_box_iou", _ptr(a), a.shape[0], _ptr(b), b.shape[0], _ptr(out), _lib.stream_ptr())
        return out

    def detectron_post(self, logits, deltas, strides, cell_anchors, image_hw, ctrness=None, score_thresh=0.05, topk=1000,
                       nms_thresh=0.5, max_det=100):
        """RetinaNet (ctrness None) / FCOS dense-detector inference. logits[l]: CUDA f32 [A*K,H,W], deltas[l]: [A*4,H,W]
        (one image) or [B,A*K,H,W] / [B,A*4,H,W] (a batch), cell_anchors[l]: A rows (x0,y0,x1,y1).
        Returns (boxes [max_det,4], scores, classes, count [1]), with a leading B on each for a batch (count [B])."""
        nl = len(logits)
        a = len(cell_anchors[0])
        batched = logits[0].dim() == 4
        b = logits[0].shape[0] if batched else 1
        lead = (b,) if batched else ()
        k = logits[0].shape[-3] // a
        keep = [t.contiguous() for t in list(logits) + list(deltas) + (list(ctrness) if ctrness is not None else [])]

        def arr(ts):
            out = (c_void_p * nl)()
            for i, t in enumerate(ts):
                out[i] = t.contiguous().data_ptr()
            return out

        dev = logits[0].device
        hs = (c_int * nl)(*[int(t.shape[-2]) for t in logits])
        ws = (c_int * nl)(*[int(t.shape[-1]) for t in logits])
        ss = (c_int * nl)(*[int(v) for v in strides])
        flat = [float(v) for lv in cell_anchors for row in lv for v in row]
        ca = (c_float * len(flat))(*flat)
        ob = torch.zeros(lead + (max_det, 4), dtype=torch.float32, device=dev)
        osc = torch.zeros(lead + (max_det,), dtype=torch.float32, device=dev)
        oc = torch.zeros(lead + (max_det,), dtype=torch.int64, device=dev)
        cnt = torch.zeros((b,), dtype=torch.int32, device=dev)
        self.ctx.call("tcb_detectron_post", b, nl, arr(keep[:nl]), arr(keep[nl:2 * nl]), arr(keep[2 * nl:]) if ctrness is not None else None,
                      hs, ws, ss, a, k, ca, 0 if ctrness is None else 1, float(score_thresh), int(topk), float(nms_thresh), int(max_det),
                      int(image_hw[0]), int(image_hw[1]), _ptr(ob), _ptr(osc), _ptr(oc), _ptr(cnt), _lib.stream_ptr())
        return ob, osc, oc, cnt

    # ---- LiDAR voxelisation + pillar decoration ------------------------------------------------------
    def voxelize(self, points, point_range, voxel_size, max_pts, max_vox, overflow="continue", f64_math=False,
                 range_mask=False, z_offset=0.0, intensity_col=-1, nfeat=None, offsets=None, out=None):
        """points: f32 CUDA [n, C]. One sweep: returns device tensors (voxels [max_vox,P,F], coords [max_vox,4] b,z,y,x,
        num_points [max_vox], m [1]); rows >= m are not meaningful. A batch: `offsets` = b+1 ascending row offsets of the
        sweeps inside `points` (host ints) and every result gains a leading batch dimension ([b,max_vox,...], m [b]; the
        batch column of coords is the sweep's index). `out` = a previous call's result tuple to write into (same shapes): a
        steady-state caller then allocates nothing per call. No host sync."""
        _need_cuda(points)
        if points.dtype != torch.float32 or points.dim() != 2 or (points.shape[0] > 0 and points.stride(1) != 1):
            raise TypeError("voxelize: points must be a float32 [n, C] tensor with contiguous rows")
        n, c = points.shape
        f = int(nfeat) if nfeat is not None else c
        dev = points.device
        b = 1 if offsets is None else len(offsets) - 1
        lead = () if offsets is None else (b,)
        if out is not None:
            voxels, coords, num, m = out
            if tuple(voxels.shape) != lead + (max_vox, max_pts, f) or m.shape[0] != b:
                raise ValueError("voxelize: `out` does not match this call's shapes")
        else:
            voxels = torch.empty(lead + (max_vox, max_pts, f), dtype=torch.float32, device=dev)
            coords = torch.empty(lead + (max_vox, 4), dtype=torch.int32, device=dev)
            num = torch.empty(lead + (max_vox,), dtype=torch.int32, device=dev)
            m = torch.empty((b,), dtype=torch.int32, device=dev)
        rng = (ctypes.c_double * 6)(*[float(v) for v in point_range])
        vs = (ctypes.c_double * 3)(*[float(v) for v in voxel_size])
        offs = None if offsets is None else (c_int * (b + 1))(*[int(v) for v in offsets])
        self.ctx.call("tcb_voxelize", _ptr(points), n, offs, b, points.stride(0) if n > 0 else c, f, rng, vs, int(max_pts), int(max_vox),
                      1 if overflow == "break" else 0, 1 if f64_math else 0, 1 if range_mask else 0, float(z_offset),
                      int(intensity_col), _ptr(voxels), _ptr(coords), _ptr(num), _ptr(m), _lib.stream_ptr())
        return voxels, coords, num, m

    def pillar_decorate(self, voxels, coords, num_points, voxel_size, point_range, m=None, out=None):
        """(M,P,F) voxels -> (M,P,F+6) decorated features, or a batch (b,M,P,F) -> (b,M,P,F+6); `m` may be the device
        count(s) from voxelize()."""
        _need_cuda(voxels, coords, num_points)
        b = voxels.shape[0] if voxels.dim() == 4 else 1
        mm, p, f = voxels.shape[-3:]
        if out is None:
            out = torch.empty(tuple(voxels.shape[:-1]) + (f + 6,), dtype=torch.float32, device=voxels.device)
        vs = (ctypes.c_double * 3)(*[float(v) for v in voxel_size])
        r0 = (ctypes.c_double * 3)(*[float(v) for v in point_range[:3]])
        self.ctx.call("tcb_pillar_decorate", _ptr(voxels), _ptr(coords), _ptr(num_points), _ptr(m) if m is not None else None,
                      b, mm, p, f, vs, r0, _ptr(out), _lib.stream_ptr())
        return out

    def dynamic_pillar(self, points, point_range, voxel_size, batch_size=1):
        """DynamicPillarVFE decoration: points f32 CUDA [n, 1+F] (batch_idx, x, y, z, ...). Returns device tensors
        (features [n, F+6], unq_inv int64 [n], voxel_coords int32 [n,4], counts int32 [2] = n_valid, n_unique)."""
        _need_cuda(points)
        pts = points.contiguous().float()
        n, ncol = pts.shape
        dev = pts.device
        feats = torch.empty((max(n, 1), ncol - 1 + 6), dtype=torch.float32, device=dev)
        inv = torch.empty((max(n, 1),), dtype=torch.int64, device=dev)
        coords = torch.empty((max(n, 1), 4), dtype=torch.int32, device=dev)
        counts = torch.zeros((2,), dtype=torch.int32, device=dev)
        rng = (ctypes.c_double * 6)(*[float(v) for v in point_range])
        vs = (ctypes.c_double * 3)(*[float(v) for v in voxel_size])
        self.ctx.call("tcb_dynamic_pillar", _ptr(pts), n, ncol, int(batch_size), rng, vs, _ptr(feats), _ptr(inv), _ptr(coords),
                      _ptr(counts), _lib.stream_ptr())
        return feats, inv, coords, counts

    # ---- 3D detection postprocess --------------------------------------------------------------------
    def rotated_iou(self, boxes_a, boxes_b):
        """boxes_iou_bev: (N,7),(M,7) f32 CUDA -> (N,M) f32."""
        _need_cuda(boxes_a, boxes_b)
        a, b = boxes_a.contiguous().float(), boxes_b.contiguous().float()
        out = torch.empty((a.shape[0], b.shape[0]), dtype=torch.float32, device=a.device)
        self.ctx.call("tcb_rotated_iou", _ptr(a), a.shape[0], _ptr(b), b.shape[0], _ptr(out), _lib.stream_ptr())
        return out

    def rotated_nms(self, boxes, scores, thresh, pre_max=4096, post_max=500, nms_type="nms_gpu"):
        """topk(pre_max) + getattr(iou3d_nms_utils, nms_type) + [:post_max] (nms_gpu: rotated, nms_normal_gpu: axis-aligned). boxes (n,7) + scores (n,) -> (keep int32 [post_max] indices into the
        input, count [1]); a batch boxes (b,n,7) + scores (b,n) -> (keep [b,post_max], count [b])."""
        _need_cuda(boxes, scores)
        bx, s = boxes.contiguous().float(), scores.contiguous().float()
        batched = bx.dim() == 3
        b, n = (bx.shape[0], bx.shape[1]) if batched else (1, bx.shape[0])
        keep = torch.empty((b, post_max) if batched else (post_max,), dtype=torch.int32, device=bx.device)
        cnt = torch.zeros((b,), dtype=torch.int32, device=bx.device)
        self.ctx.call("tcb_rotated_nms", _ptr(bx), _ptr(s), b, n, float(thresh), _nms_type(nms_type), int(pre_max), int(post_max), _ptr(keep), _ptr(cnt),
                      _lib.stream_ptr())
        return keep, cnt

    def multi_classes_nms(self, cls_scores, box_preds, nms_thresh, pre_max=4096, post_max=500, score_thresh=None, nms_type="nms_gpu"):
        """model_nms_utils.multi_classes_nms (OPCD/pcdet/models/model_utils/model_nms_utils.py:28-66): per class the score cut,
        top pre_max, NMS of NMS_TYPE and the first post_max — the classes are the batch of one tcb_rotated_nms call (scores
        under the cut enter as -inf: they order last, cannot suppress a row that counts, and are dropped from the kept lists).
        cls_scores (N, C), box_preds (N, 7 + extra) CUDA -> (pred_scores, pred_labels int64, pred_boxes), classes in order."""
        _need_cuda(cls_scores, box_preds)
        sc = cls_scores.float().t().contiguous()                                  # [C, N]
        c, n = sc.shape
        bp = box_preds.float().contiguous()
        if n == 0 or c == 0:
            return sc.new_zeros((0,)), torch.zeros((0,), dtype=torch.int64, device=sc.device), bp[:0]
        if score_thresh is not None:
            sc = torch.where(sc >= float(score_thresh), sc, torch.full_like(sc, float("-inf")))
        boxes7 = bp[:, :7].unsqueeze(0).expand(c, n, 7).contiguous()
        keep, cnt = self.rotated_nms(boxes7, sc, nms_thresh, pre_max, post_max, nms_type=nms_type)
        kept_scores = torch.gather(sc, 1, keep.long().clamp_(0, n - 1))          # [C, post_max]
        valid = (torch.arange(keep.shape[1], device=sc.device)[None, :] < cnt[:, None]) & (kept_scores > float("-inf"))
        labels = torch.arange(c, device=sc.device, dtype=torch.int64)[:, None].expand_as(keep)
        rows = keep.long()[valid]                                                 # class-major, kept order inside a class
        return kept_scores[valid], labels[valid], bp[rows]

    def pp_anchor_post(self, cls_preds, box_preds, dir_preds, anchor_range, anchor_sizes, anchor_bottoms, anchor_rotations,
                       dir_offset=0.78539, dir_limit_offset=0.0, num_dir_bins=2, score_thresh=0.1, nms_thresh=0.01,
                       pre_max=4096, post_max=500, out=None, nms_type="nms_gpu"):
        """PointPillars head outputs of one frame — cls (H,W,A*nc), box (H,W,A*7), dir (H,W,A*bins) or None — or of a
        batch (B,H,W,...). Returns device tensors (boxes [post_max,7], scores [post_max], labels int64 [post_max],
        count [1]), with a leading B on each for a batch (count [B])."""
        _need_cuda(cls_preds, box_preds)
        batched = cls_preds.dim() == 4
        b = cls_preds.shape[0] if batched else 1
        lead = (b,) if batched else ()
        h, w = cls_preds.shape[-3], cls_preds.shape[-2]
        nc = len(anchor_sizes)
        nrot = len(anchor_rotations[0])
        dev = cls_preds.device
        if out is not None:
            ob, osc, ol, cnt = out            # a previous call's result tuple (same shapes): rows beyond count keep older values
        else:
            ob = torch.zeros(lead + (post_max, 7), dtype=torch.float32, device=dev)
            osc = torch.zeros(lead + (post_max,), dtype=torch.float32, device=dev)
            ol = torch.zeros(lead + (post_max,), dtype=torch.int64, device=dev)
            cnt = torch.zeros((b,), dtype=torch.int32, device=dev)
        rng = (c_float * 6)(*[float(np.float32(v)) for v in anchor_range])
        sizes = (c_float * (nc * 3))(*[float(v) for s in anchor_sizes for v in s])
        bots = (c_float * nc)(*[float(v) for v in anchor_bottoms])
        rots = (c_float * (nc * nrot))(*[float(v) for r in anchor_rotations for v in r])
        self.ctx.call("tcb_pp_anchor_post", _ptr(cls_preds.contiguous()), _ptr(box_preds.contiguous()),
                      _ptr(dir_preds.contiguous()) if dir_preds is not None else None, b, h, w, nc, nrot, rng, sizes, bots, rots,
                      float(dir_offset), float(dir_limit_offset), int(num_dir_bins), float(score_thresh), float(nms_thresh),
                      _nms_type(nms_type), int(pre_max), int(post_max), _ptr(ob), _ptr(osc), _ptr(ol), _ptr(cnt), _lib.stream_ptr())
        return ob, osc, ol, cnt

    def centerpoint_post(self, heads, class_id_map, feature_map_stride, voxel_size, point_range, limit_range, k=500,
                         score_thresh=0.1, nms_thresh=0.2, pre_max=1000, post_max=83, out=None, nms_type="nms_gpu"):
        """heads: list of dicts with CUDA f32 tensors hm [c,H,W], center [2,H,W], center_z [1,H,W], dim [3,H,W],
        rot [2,H,W], optional vel [2,H,W] — or each with a leading batch dimension [B,...]; class_id_map: list (per head)
        of global 0-based class ids. Returns (boxes [heads*post_max,9], scores, labels int64 1-based, count [1]), with a
        leading B on each for a batch (count [B])."""
        nh = len(heads)
        keep_alive = []

        def ptrs(name, optional=False):
            arr = (c_void_p * nh)()
            for i, hd in enumerate(heads):
                t = hd.get(name)
                if t is None:
                    if not optional:
                        raise KeyError(name)
                    arr[i] = None
                else:
                    t = t.contiguous()
                    keep_alive.append(t)
                    arr[i] = t.data_ptr()
            return arr

        hm = heads[0]["hm"]
        _need_cuda(hm)
        batched = hm.dim() == 4
        b = hm.shape[0] if batched else 1
        lead = (b,) if batched else ()
        h, w = hm.shape[-2], hm.shape[-1]
        dev = hm.device
        ncls = (c_int * nh)(*[int(hd["hm"].shape[-3]) for hd in heads])
        cmap = [int(v) for m in class_id_map for v in m]
        cmap_arr = (c_int * len(cmap))(*cmap)
        if out is not None:
            ob, osc, ol, cnt = out            # a previous call's result tuple (same shapes): rows beyond count keep older values
        else:
            ob = torch.zeros(lead + (nh * post_max, 9), dtype=torch.float32, device=dev)
            osc = torch.zeros(lead + (nh * post_max,), dtype=torch.float32, device=dev)
            ol = torch.zeros(lead + (nh * post_max,), dtype=torch.int64, device=dev)
            cnt = torch.zeros((b,), dtype=torch.int32, device=dev)
        has_vel = all(hd.get("vel") is not None for hd in heads)
        self.ctx.call("tcb_centerpoint_post", b, nh, ptrs("hm"), ptrs("center"), ptrs("center_z"), ptrs("dim"), ptrs("rot"),
                      ptrs("vel", True) if has_vel else None, ncls, cmap_arr, h, w, int(feature_map_stride),
                      (c_float * 2)(float(voxel_size[0]), float(voxel_size[1])), (c_float * 2)(float(point_range[0]), float(point_range[1])),
                      (c_float * 6)(*[float(v) for v in limit_range]), int(k), float(score_thresh), float(nms_thresh),
                      _nms_type(nms_type), int(pre_max), int(post_max), _ptr(ob), _ptr(osc), _ptr(ol), _ptr(cnt), _lib.stream_ptr())
        return ob, osc, ol, cnt

    def class_score_filter(self, labels, scores, groups):
        """groups: list of (class, threshold) in output order. Returns (indices int32 [len(groups)*n], count [1])."""
        _need_cuda(labels, scores)
        n = labels.shape[0]
        lab, sc = labels.contiguous().to(torch.int64), scores.contiguous().float()
        out = torch.empty((max(1, len(groups) * n),), dtype=torch.int32, device=lab.device)
        cnt = torch.zeros((1,), dtype=torch.int32, device=lab.device)
        gc = (c_int * max(1, len(groups)))(*[int(g[0]) for g in groups])
        gt = (c_float * max(1, len(groups)))(*[float(g[1]) for g in groups])
        self.ctx.call("tcb_class_score_filter", _ptr(lab), _ptr(sc), n, len(groups), gc, gt, _ptr(out), _ptr(cnt), _lib.stream_ptr())
        return out, cnt

    # ---- after the path: box rescale + overlay -------------------------------------
    def scale_boxes(self, det, orig_size, input_size, normalized=False, math_mode=0):
        """_scale_box_array (evaluate_inference.py:237-256) on (n,6) float64 CUDA rows; math_mode 1/2 = _scale_boxes
        (ros_inference.py:100-115) under numpy < 2 / >= 2."""
        _need_cuda(det)
        d = det.contiguous().to(torch.float64).reshape(-1, 6)
        out = torch.empty_like(d)
        self.ctx.call("tcb_scale_boxes", _ptr(d), d.shape[0], int(orig_size[0]), int(orig_size[1]), int(input_size[0]), int(input_size[1]),
                      int(bool(normalized)), int(math_mode), _ptr(out), _lib.stream_ptr())
        return out

    def draw_detections(self, frames, det, counts, class_names, color, input_size, thickness=2, labels=True, math_mode=1,
                        normalized=False, per_class_color=False, font_scale=0.5, text_thickness=2, text_dy=-10):
        """The drawing loop of RosInference._callback (ros_inference.py:148-169) in place on CUDA frames
        [b,h,w,3] (or [h,w,3]) uint8; det [b,max_det,6] float64 + counts [b] int32 as extract_boxes_device returns.
        labels: False / True ('{:.2f} {name}') / "name" (the class name alone, as plot_boxes_cv2 writes it); the text is
        cv2.putText(text, (x1, int(y1 + text_dy)), FONT_HERSHEY_SIMPLEX, font_scale, color, text_thickness)."""
        _need_cuda(frames, det)
        if frames.dtype != torch.uint8 or frames.shape[-1] != 3 or frames.dim() not in (3, 4):
            raise TypeError("draw_detections: frames must be uint8 [b,h,w,3] or [h,w,3]")
        f4 = frames if frames.dim() == 4 else frames.unsqueeze(0)
        if f4.stride(3) != 1 or f4.stride(2) != 3:
            raise TypeError("draw_detections: pixels must be packed (stride 3,1); rows and frames may be strided")
        b, h, w = f4.shape[:3]
        d = det.contiguous().to(torch.float64).reshape(b, -1, 6)
        cnt = None if counts is None else counts.contiguous().to(torch.int32)
        names = b"".join(str(n).encode("latin-1", "replace") + b"\0" for n in class_names)
        cols = bytes(int(v) & 255 for c in (color if per_class_color else [color]) for v in c)
        self.ctx.call("tcb_draw_detections", _ptr(f4), b, h, w, f4.stride(1), f4.stride(0) if b > 1 else h * f4.stride(1), _ptr(d),
                      _ptr(cnt) if cnt is not None else None, d.shape[1], int(input_size[0]), int(input_size[1]), int(bool(normalized)),
                      int(math_mode), names, len(class_names), cols, int(bool(per_class_color)), int(thickness),
                      2 if labels == "name" else int(bool(labels)), ctypes.c_double(float(font_scale)), int(text_thickness), int(text_dy),
                      _lib.stream_ptr())
        return frames

    # ---- ROS message decode/encode around the 3D path ------------------------------
    def pointcloud2_unpack(self, data, width, height, point_step, row_step, offsets, datatypes, is_bigendian=False, skip_nans=True):
        """read_points(...) of ros_inference3d.py:125 on the device. data: CUDA uint8 tensor with the message bytes; offsets /
        datatypes of the selected fields in ascending offset. Returns (points f32 [width*height, k] with sentinel rows
        after the kept ones, count int32 [1])."""
        _need_cuda(data)
        d = data.contiguous().view(torch.uint8).reshape(-1)
        n, k = int(width) * int(height), len(offsets)
        need = (int(height) - 1) * int(row_step) + int(width) * int(point_step) if n else 0
        if d.numel() < need:
            raise ValueError(f"pointcloud2_unpack: message data holds {d.numel()} bytes, the layout needs {need}")
        out = torch.empty((n, k), dtype=torch.float32, device=d.device)
        cnt = torch.zeros((1,), dtype=torch.int32, device=d.device)
        offs, dts = (c_int * max(1, k))(*[int(v) for v in offsets]), (c_int * max(1, k))(*[int(v) for v in datatypes])
        self.ctx.call("tcb_pointcloud2_unpack", _ptr(d) if n else None, int(width), int(height), int(point_step), int(row_step), k, offs, dts,
                      int(bool(is_bigendian)), int(bool(skip_nans)), _ptr(out) if n else None, _ptr(cnt), _lib.stream_ptr())
        return out, cnt

    def boxes_to_jsk(self, boxes, scores, labels, label_sel=2, score_thresh=0.5, z_offset=1.5):
        """BoundingBoxArray fields (ros_inference3d.py:150-177). Returns (rows f64 [n,12], indices int32 [n], count int32 [1])."""
        _need_cuda(boxes, scores, labels)
        b, s, l = boxes.contiguous().float().reshape(-1, 7), scores.contiguous().float(), labels.contiguous().to(torch.int64)
        n = b.shape[0]
        out = torch.zeros((max(1, n), 12), dtype=torch.float64, device=b.device)
        idx = torch.zeros((max(1, n),), dtype=torch.int32, device=b.device)
        cnt = torch.zeros((1,), dtype=torch.int32, device=b.device)
        self.ctx.call("tcb_boxes_to_jsk", _ptr(b) if n else None, _ptr(s) if n else None, _ptr(l) if n else None, n, int(label_sel),
                      float(score_thresh), float(z_offset), _ptr(out), _ptr(idx), _ptr(cnt), _lib.stream_ptr())
        return out, idx, cnt


    # ---- evaluation math (§8f rank 4) --------------------------------------------------------------
    @staticmethod
    def _offsets(counts):
        off = (c_int * (len(counts) + 1))()
        for i, c in enumerate(counts):
            off[i + 1] = off[i] + int(c)
        return off

    def _eval_outputs(self, frames, total_preds, max_classes, nt, dev):
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)
        return dict(correct=z((max(1, total_preds), 10), torch.uint8), matched=z((frames,), torch.int32),
                    num_classes=z((frames,), torch.int32), classes=z((frames, max_classes), torch.int32),
                    p=z((frames, max_classes), torch.float64), r=z((frames, max_classes), torch.float64),
                    f1=z((frames, max_classes), torch.float64), ap=z((frames, max_classes, nt), torch.float64))

    def eval_match(self, gts, gt_counts, preds, pred_counts):
        """calculate_metrics' matching (evaluate_inference.py:408-425) for a batch of frames. gts / preds: CUDA float64
        [total,6] rows [x1,y1,x2,y2,score,class], frame f owning `counts[f]` consecutive rows.
        Returns (correct u8 [total_preds,10], matched i32 [frames])."""
        _need_cuda(gts, preds)
        g, p = gts.contiguous().double().reshape(-1, 6), preds.contiguous().double().reshape(-1, 6)
        f = len(gt_counts)
        o = self._eval_outputs(f, p.shape[0], 1, 10, g.device)
        self.ctx.call("tcb_eval_match", _ptr(g), self._offsets(gt_counts), _ptr(p), self._offsets(pred_counts), f,
                      _ptr(o["correct"]), _ptr(o["matched"]), _lib.stream_ptr())
        return o["correct"][:p.shape[0]], o["matched"]

    def ap_per_class(self, tp, conf, pred_cls, pred_counts, target_cls, target_counts, max_classes=16):
        """ap_per_class (evaluate_inference.py:158-218) per frame: tp CUDA u8/bool [total,nt], conf / pred_cls float64
        [total], target_cls float64 [total_targets]. Returns dict(num_classes, classes, p, r, f1, ap) of CUDA tensors."""
        _need_cuda(tp, conf, pred_cls, target_cls)
        t = tp.contiguous().to(torch.uint8)
        c, pc, tc = conf.contiguous().double(), pred_cls.contiguous().double(), target_cls.contiguous().double()
        f, nt = len(pred_counts), int(t.shape[1])
        o = self._eval_outputs(f, 0, max_classes, nt, t.device)
        self.ctx.call("tcb_ap_per_class", _ptr(t), nt, _ptr(c), _ptr(pc), 1, self._offsets(pred_counts), _ptr(tc), 1,
                      self._offsets(target_counts), f, int(max_classes), _ptr(o["num_classes"]), _ptr(o["classes"]),
                      _ptr(o["p"]), _ptr(o["r"]), _ptr(o["f1"]), _ptr(o["ap"]), _lib.stream_ptr())
        return {k: o[k] for k in ("num_classes", "classes", "p", "r", "f1", "ap")}

    def calculate_metrics(self, gts, gt_counts, preds, pred_counts, max_classes=16):
        """Matching + per-frame ap_per_class in two launches (evaluate_inference.py:400-446). Returns the dict of
        `_eval_outputs`; frames with matched == 0 have no statistics (the reference skips them)."""
        _need_cuda(gts, preds)
        g, p = gts.contiguous().double().reshape(-1, 6), preds.contiguous().double().reshape(-1, 6)
        f = len(gt_counts)
        o = self._eval_outputs(f, p.shape[0], max_classes, 10, g.device)
        self.ctx.call("tcb_calculate_metrics", _ptr(g), self._offsets(gt_counts), _ptr(p), self._offsets(pred_counts), f,
                      int(max_classes), _ptr(o["correct"]), _ptr(o["matched"]), _ptr(o["num_classes"]), _ptr(o["classes"]),
                      _ptr(o["p"]), _ptr(o["r"]), _ptr(o["f1"]), _ptr(o["ap"]), _lib.stream_ptr())
        o["correct"] = o["correct"][:p.shape[0]]
        return o


def yolo_nms_static(pred, conf_thres=0.6, iou_thres=0.45, classes=None, agnostic=False, max_det=300,
                    max_nms=30000, out=None, counts=None):
    """tcb_yolo_nms_static: the YOLOv5 postprocess for callers that may not allocate GPU memory (a plugin
    in a host process that owns all of HBM, a frame callback that cannot take cudaMalloc's device-wide
    sync). Needs no Engine/context; scratch is a fixed ~8.4 MB device array inside the library, so sizes
    are capped (b <= 64, b*n <= 131072) and overlapping calls race — keep them on one stream. Pass `out`
    ([b,max_det,6] f64) and `counts` ([2,b] i32) to keep this wrapper allocation-free as well."""
    _need_cuda(pred)
    if pred.dtype != torch.float32 or pred.dim() != 3 or not pred.is_contiguous():
        raise TypeError("yolo_nms_static: pred must be a contiguous float32 [b,n,5+nc] tensor")
    b, n, c = pred.shape
    if out is None:
        out = torch.zeros((b, max_det, 6), dtype=torch.float64, device=pred.device)
    if counts is None:
        counts = torch.zeros((2, b), dtype=torch.int32, device=pred.device)
    cls_arr, ncls = None, 0
    if classes is not None:
        ncls = len(classes)
        cls_arr = (c_int * max(ncls, 1))(*[int(v) for v in classes])
    l = _lib.lib()
    with torch.cuda.device(pred.device):
        rc = l.tcb_yolo_nms_static(_ptr(pred), b, n, c - 5, float(conf_thres), float(iou_thres), cls_arr, ncls,
                                   1 if agnostic else 0, int(max_det), int(max_nms), _ptr(out), _ptr(counts[0]),
                                   _ptr(counts[1]), _lib.stream_ptr())
    if rc != 0:
        raise _lib.TcbError(rc, (l.tcb_last_error(None) or b"").decode())
    return out, counts[0], counts[1]
