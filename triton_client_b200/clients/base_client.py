"""Mirror of clients/base_client.py:5-104 — the Client ABC and the Triton-metadata check. Kept verbatim
in behaviour (SURVEY §8b "boundary"); not accelerated. tritonclient is optional here: only the two
ModelInput format constants are needed."""
from abc import ABC, abstractmethod

try:  # pragma: no cover - tritonclient is not in the build image
    import tritonclient.grpc.model_config_pb2 as mc
    FORMAT_NHWC, FORMAT_NCHW = mc.ModelInput.FORMAT_NHWC, mc.ModelInput.FORMAT_NCHW
except Exception:
    mc = None
    FORMAT_NHWC, FORMAT_NCHW = 1, 2  # model_config.proto: FORMAT_NONE=0, FORMAT_NHWC=1, FORMAT_NCHW=2


class Client(ABC):
    """Declares the functionality for several clients for triton server"""

    def __init__(self, engine=None):
        self._clients = {}
        self._engine = engine

    @abstractmethod
    def register_client(self, clienttype, client):
        """Implement the method to register the client for"""

    @abstractmethod
    def get_postprocess(self):
        """ """

    @abstractmethod
    def get_preprocess(self):
        """ """

    def parse_model(self, model_metadata, model_config):
        """Image models: exactly one input with 3 dims in NCHW or NHWC; returns
        (input name, output names, c, h, w, format, datatype)."""
        if len(model_metadata.inputs) != 1:
            raise Exception("expecting 1 input, got {}".format(len(model_metadata.inputs)))
        if len(model_config.input) != 1:
            raise Exception("expecting 1 input in model configuration, got {}".format(len(model_config.input)))
        input_metadata = model_metadata.inputs[0]
        input_config = model_config.input[0]
        output_metadata = [output for output in model_metadata.outputs]
        expected_input_dims = 3
        if len(input_metadata.shape) != expected_input_dims:
            raise Exception("expecting input to have {} dimensions, model '{}' input has {}".format(
                expected_input_dims, model_metadata.name, len(input_metadata.shape)))
        if input_config.format != FORMAT_NCHW and input_config.format != FORMAT_NHWC:
            raise Exception("unexpected input format {}, expecting FORMAT_NCHW or FORMAT_NHWC".format(input_config.format))
        if input_config.format == FORMAT_NHWC:
            h, w, c = input_metadata.shape[0], input_metadata.shape[1], input_metadata.shape[2]
        else:
            c, h, w = input_metadata.shape[0], input_metadata.shape[1], input_metadata.shape[2]
        return (input_metadata.name, [output.name for output in output_metadata], c, h, w,
                input_config.format, input_metadata.datatype)
