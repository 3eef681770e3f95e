from .base_postprocess import Postprocess
from .yolov5_postprocess import Yolov5postprocess
from .detectron_postprocess import FCOSpostprocess
