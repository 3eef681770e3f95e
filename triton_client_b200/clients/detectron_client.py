"""Mirror of clients/detectron_client.py:4-21."""
from .base_client import Client
from .preprocess import detectron_preprocess
from .postprocess import detectron_postprocess


class FCOS_client(Client):
    def __init__(self, engine=None):
        super().__init__(engine)

    def register_client(self, clienttype, client):
        self._clients[clienttype] = client

    def get_preprocess(self):
        return detectron_preprocess.FCOSpreprocess(self._engine)

    def get_postprocess(self):
        return detectron_postprocess.FCOSpostprocess(self._engine)
