"""Drop-in for clients/preprocess/detectron_preprocess.py:12-24 — same as the YOLOv5 one without /255
(detectron2 normalises inside the served model)."""
from .yolov5_preprocess import _ImagePreprocess


class FCOSpreprocess(_ImagePreprocess):
    DIV = 1.0
