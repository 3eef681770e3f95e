"""Drop-in for clients/postprocess/yolov5_postprocess.py: same class, same `extract_boxes` signature,
same return value (list of (n,6) float64 CPU tensors, or the caught exception object), with filter,
decode, ordering and NMS done on the device by tcb_yolo_nms."""
import os

import numpy as np
import torch

from .base_postprocess import Postprocess

_DATA = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "data")


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
