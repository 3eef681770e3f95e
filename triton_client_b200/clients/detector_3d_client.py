"""Mirror of clients/detector_3d_client.py:8-91 (Pointpillars_client)."""
from .base_client import Client
from .preprocess.preprocess_3d import PointpillarPreprocess
from .preprocess.voxelize import det3DPreprocess  # noqa: F401
from .postprocess.detector_3d_postprocess import PointPillarPostprocess


class Pointpillars_client(Client):
    def __init__(self, engine=None):
        super().__init__(engine)

    def register_client(self, clienttype, client):
        self._clients[clienttype] = client

    def get_preprocess(self):
        return PointpillarPreprocess(self._engine)

    def get_postprocess(self):
        return PointPillarPostprocess(self._engine)

    def parse_model(self, model_metadata, model_config):
        if len(model_metadata.inputs) != 3:     # voxels, coords, numpoints
            raise Exception("expecting 3 input, got {}".format(len(model_metadata.inputs)))
        if len(model_metadata.outputs) != 3:    # bbox_preds, dir_scores, scores
            raise Exception("expecting 3 output, got {}".format(len(model_metadata.outputs)))
        input_metadata = [{'name': i.name, 'shape': i.shape, 'dtype': i.datatype} for i in model_metadata.inputs]
        output_metadata = [{'name': o.name, 'shape': o.shape, 'dtype': o.datatype} for o in model_metadata.outputs]
        return (input_metadata, output_metadata)
