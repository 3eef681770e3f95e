"""Mirror of clients/postprocess/base_postprocess.py: the helpers every postprocess object inherits.
Wire parsing is a reinterpretation, not a per-element loop; the numeric helpers run on the GPU."""
from abc import ABC, abstractmethod

import numpy as np
import torch

from .. import wire


class Postprocess(ABC):
    def __init__(self, engine=None):
        self._engine = engine
        self._stager = None

    @property
    def engine(self):
        if self._engine is None:
            self._engine = wire.default_engine()
        return self._engine

    @property
    def stager(self):
        if self._stager is None:
            self._stager = wire.Stager(self.engine.device)
        return self._stager

    # base_postprocess.py:15-25 — same values and dtype (object array of Python floats), no struct loop
    def deserialize_bytes_float(self, encoded_tensor):
        return np.frombuffer(encoded_tensor, dtype="<f4").astype(np.float64).astype(np.object_)

    # base_postprocess.py:27-37 — 'l' (8 bytes, native) as in clients/; utils/postprocess.py uses 'i'
    def deserialize_bytes_int(self, encoded_tensor):
        return np.frombuffer(encoded_tensor, dtype="<i8").astype(np.object_)

    # base_postprocess.py:39-46
    def xywh2xyxy(self, x):
        y = x.clone() if isinstance(x, torch.Tensor) else np.copy(x)
        y[:, 0] = x[:, 0] - x[:, 2] / 2
        y[:, 1] = x[:, 1] - x[:, 3] / 2
        y[:, 2] = x[:, 0] + x[:, 2] / 2
        y[:, 3] = x[:, 1] + x[:, 3] / 2
        return y

    # base_postprocess.py:48-70 — pairwise IoU, on the GPU (tcb_box_iou); returns a CPU tensor like the reference
    def box_iou(self, box1, box2):
        dev = self.engine.device
        a = torch.as_tensor(box1, dtype=torch.float32).to(dev)
        b = torch.as_tensor(box2, dtype=torch.float32).to(dev)
        return self.engine.box_iou(a, b).cpu()

    # base_postprocess.py:72-106 — numpy greedy NMS, on the GPU (tcb_nms_axis_aligned); returns np.array of kept indices
    def nms_cpu(self, boxes, confs, nms_thresh=0.5, min_mode=False):
        dev = self.engine.device
        b = torch.as_tensor(np.asarray(boxes, dtype=np.float32)).to(dev)
        s = torch.as_tensor(np.asarray(confs, dtype=np.float32)).to(dev)
        keep, cnt = self.engine.nms_axis_aligned(b, s, nms_thresh, min_mode)
        kept, cut = (int(v) for v in cnt.cpu())
        if cut or (kept >= keep.shape[0] and b.shape[0] > kept):
            # the reference has no bound; the kernel considers the 4096 best boxes and keeps at most 1024
            raise OverflowError("nms_cpu: %d boxes (limit 4096) / %d kept (limit %d)" % (b.shape[0], kept, keep.shape[0]))
        return keep[:kept].cpu().numpy().astype(np.int64)

    @abstractmethod
    def postprocess(self):
        """ """
