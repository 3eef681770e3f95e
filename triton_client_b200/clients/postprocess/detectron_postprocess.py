"""Drop-in for clients/postprocess/detectron_postprocess.py:26-37 (deserialise boxes/classes/scores)."""
import numpy as np

from .base_postprocess import Postprocess


class FCOSpostprocess(Postprocess):
    def __init__(self, engine=None):
        super().__init__(engine)

    def postprocess(self):
        pass

    def load_class_names(self, namesfile='./data/crop.names'):
        with open(namesfile, 'r') as fp:
            return [line.rstrip() for line in fp.readlines()]

    def extract_boxes(self, prediction):
        boxes = self.deserialize_bytes_float(prediction.raw_output_contents[0])
        boxes = np.reshape(boxes, prediction.outputs[0].shape)
        class_ids = self.deserialize_bytes_int(prediction.raw_output_contents[1])
        scores = self.deserialize_bytes_float(prediction.raw_output_contents[2])
        scores = np.reshape(scores, prediction.outputs[2].shape)
        return boxes, class_ids, scores
