"""Drop-in for clients/postprocess/yolov5_postprocess.py: same class, same `extract_boxes` signature,
same return value (list of (n,6) float64 CPU tensors, or the caught exception object), with filter,
decode, ordering and NMS done on the device by tcb_yolo_nms."""
import os

import numpy as np
import torch

from .base_postprocess import Postprocess

_DATA = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "data")


class PendingBoxes:
    """Handle returned by Yolov5postprocess.extract_boxes_async: `.result()` waits for the queued work and returns what
    extract_boxes returns (the list of (n,6) tensors, or the exception object)."""

    def __init__(self, finish=None, value=None, keep=()):
        self._finish, self._value, self._keep = finish, value, keep

    def result(self):
        if self._finish is not None:
            try:
                self._value = self._finish()
            except Exception as e:     # same convention as extract_boxes: errors are returned
                self._value = e
            self._finish, self._keep = None, ()
        return self._value


class Yolov5postprocess(Postprocess):
    def __init__(self, engine=None):
        super().__init__(engine)

    def postprocess(self):
        pass

    def load_class_names(self, namesfile='./data/coco.names'):
        class_names = []
        with open(namesfile, 'r') as fp:
            lines = fp.readlines()
        for line in lines:
            class_names.append(line.rstrip())
        return class_names

    def extract_boxes(self, prediction, conf_thres=0.6, iou_thres=0.45, classes=None, agnostic=False,
                      multi_label=False, labels=(), max_det=300):
        """Runs Non-Maximum Suppression (NMS) on inference results

        Returns:
             list of detections, on (n,6) tensor per image [xyxy, conf, cls]
        Errors are returned, not raised (yolov5_postprocess.py:43,124-125).
        """
        try:
            shape = tuple(int(v) for v in prediction.outputs[0].shape)
            pred = self.stager.to_device(prediction.raw_output_contents[0], np.float32, shape, "yolo_in")
            out, counts = self.extract_boxes_device(pred, conf_thres, iou_thres, classes, agnostic, multi_label, labels, max_det)
            rows = self.stager.to_host(out, "yolo_out", copy=False)
            output = [torch.zeros((0, 6))] * shape[0]
            for xi in range(shape[0]):
                n = int(counts[xi])
                if n:
                    output[xi] = torch.from_numpy(rows[xi, :n].copy())
            return output
        except Exception as e:
            return e

    def extract_boxes_async(self, prediction, conf_thres=0.6, iou_thres=0.45, classes=None, agnostic=False,
                            multi_label=False, labels=(), max_det=300):
        """extract_boxes for a pipelined caller (no counterpart in the reference, whose loop is one frame at a time):
        the upload of the response, the kernels and the download of the <= max_det rows are queued on this object's own
        stream and the call returns at once, so the caller can run the next batch's preprocess meanwhile (the response's
        upload then shares the PCIe link with the preprocess download instead of waiting for it). `.result()` on the
        returned handle gives exactly what extract_boxes returns. The response must stay unchanged until then; at most two
        calls in flight."""
        try:
            if not (0 <= conf_thres <= 1 and 0 <= iou_thres <= 1) or multi_label or (labels and any(len(l) for l in labels)):
                return PendingBoxes(value=self.extract_boxes(prediction, conf_thres, iou_thres, classes, agnostic, multi_label, labels, max_det))
            dev = self.engine.device
            if getattr(self, "_side", None) is None:
                self._side = torch.cuda.Stream(dev)
            side = self._side
            side.wait_stream(torch.cuda.current_stream(dev))
            shape = tuple(int(v) for v in prediction.outputs[0].shape)
            self._turn = turn = (getattr(self, "_turn", 0) + 1) & 1     # two sets of host buffers: at most two calls in flight
            with torch.cuda.stream(side):
                pred = self.stager.to_device(prediction.raw_output_contents[0], np.float32, shape, f"yolo_in_async{turn}")
                out, cnt, xc = self.engine.yolo_nms(pred, conf_thres=conf_thres, iou_thres=iou_thres, classes=classes,
                                                    agnostic=agnostic, max_det=max_det)
                both = torch.stack((cnt, xc))
                hb = self.stager.pinned(f"yolo_counts{turn}", both.numel() * 4).view(torch.int32).reshape(both.shape)
                hb.copy_(both, non_blocking=True)
                hr = self.stager.pinned(f"yolo_out_async{turn}", out.numel() * 8).view(torch.float64).reshape(out.shape)
                hr.copy_(out, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(side)

            def finish():
                ev.synchronize()
                counts, rows = hb.numpy(), hr.numpy()
                assert counts[1].sum() > 0, 'No valid detections found'
                output = [torch.zeros((0, 6))] * shape[0]
                for xi in range(shape[0]):
                    n = int(counts[0, xi])
                    if n:
                        output[xi] = torch.from_numpy(rows[xi, :n].copy())
                return output
            return PendingBoxes(finish=finish, keep=(pred, out, both, prediction))
        except Exception as e:
            return PendingBoxes(value=e)

    def extract_boxes_device(self, pred, conf_thres=0.6, iou_thres=0.45, classes=None, agnostic=False,
                             multi_label=False, labels=(), max_det=300):
        """Device-resident variant: pred is a CUDA f32 [B,N,5+nc] tensor; returns (out f64 [B,max_det,6]
        on the device, counts as a host int array). Raises what the reference's body would raise."""
        nc = pred.shape[2] - 5
        if not (0 <= conf_thres <= 1 and 0 <= iou_thres <= 1):
            # error path, in the reference's order (:44-46): any(xc) is asserted before the range checks
            assert bool((pred[..., 4].double() > conf_thres).any().item()), 'No valid detections found'
            assert 0 <= conf_thres <= 1, f'Invalid Confidence threshold {conf_thres}, valid values are between 0.0 and 1.0'
            assert 0 <= iou_thres <= 1, f'Invalid IoU {iou_thres}, valid values are between 0.0 and 1.0'
        out, cnt, xc = self.engine.yolo_nms(pred, conf_thres=conf_thres, iou_thres=iou_thres, classes=classes,
                                            agnostic=agnostic, max_det=max_det)
        both = torch.stack((cnt, xc)).cpu().numpy()
        assert both[1].sum() > 0, 'No valid detections found'
        multi_label &= nc > 1
        if multi_label:  # the reference calls ndarray.nonzero(as_tuple=False) here (:83) and returns that TypeError
            raise TypeError("nonzero() takes no keyword arguments")
        if labels and any(len(l) for l in labels):  # :64-71 builds torch tensors from a numpy array's .device
            raise AttributeError("'numpy.ndarray' object has no attribute 'device'")
        return out, both[0]

    def plot_boxes_cv2(self, img, boxes, savename=None, class_names=None, color=None):
        """clients/postprocess/yolov5_postprocess.py:127-169, drawn on the GPU with OpenCV's own raster rules
        (clients/overlay.py::plot_boxes_cv2): normalised boxes, per-class colours, 1.2-scale class-name labels."""
        from ..overlay import plot_boxes_cv2
        return plot_boxes_cv2(img, boxes, savename, class_names, color, engine=self.engine)
