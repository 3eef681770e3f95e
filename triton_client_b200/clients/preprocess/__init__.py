from .base_preprocess import Preprocess
from .yolov5_preprocess import Yolov5preprocess
from .detectron_preprocess import FCOSpreprocess
from .voxelize import det3DPreprocess
from .preprocess_3d import PointpillarPreprocess
