"""Mirror of clients/yolov5_client.py:6-23."""
from .base_client import Client
from .preprocess import yolov5_preprocess
from .postprocess import yolov5_postprocess


class Yolov5client(Client):
    def __init__(self, engine=None):
        super().__init__(engine)

    def register_client(self, clienttype, client):
        self._clients[clienttype] = client

    def get_preprocess(self):
        return yolov5_preprocess.Yolov5preprocess(self._engine)

    def get_postprocess(self):
        return yolov5_postprocess.Yolov5postprocess(self._engine)
