"""Drop-in for the reference's `clients` package (clients/__init__.py:1-9): same names, same module
paths below this package, GPU work behind them. See INTEGRATION.md for how main.py/main3d.py pick it up."""
from .yolov5_client import Yolov5client
from .base_client import Client
from .detectron_client import FCOS_client
from .postprocess import *  # noqa: F401,F403
from .preprocess import *  # noqa: F401,F403
from .detector_3d_client import Pointpillars_client
from .wire import HostResponse

# north_star spellings
YOLOv5Client = Yolov5client
DetectronClient = FCOS_client
Detector3DClient = Pointpillars_client
