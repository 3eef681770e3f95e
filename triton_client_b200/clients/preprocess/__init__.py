from .base_preprocess import Preprocess
from .yolov5_preprocess import Yolov5preprocess
from .detectron_preprocess import FCOSpreprocess
