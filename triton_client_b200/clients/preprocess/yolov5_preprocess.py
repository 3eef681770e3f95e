"""Drop-in for clients/preprocess/yolov5_preprocess.py (same class, same method, same results), with the
work done by the fused sm_100a preprocess kernel through the C ABI. Added methods (never signature
changes): fused resize+adjust and batched/device variants."""
import numpy as np
import torch

from .base_preprocess import Preprocess
from .. import wire


class _ImagePreprocess(Preprocess):
    DIV = 255.0  # yolov5_preprocess.py:22 `cv_image /= 255.0`

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

    def preprocess(self):
        pass

    # ---- reference surface -----------------------------------------------------------------------
    def image_adjust(self, cv_image):
        '''
        cv_image: input image in RGB order (HxWx3 uint8, already at model size)
        return: Normalized image in BCHW dimensions (1,3,H,W) float32, host ndarray.
        The input is not modified (the caller keeps drawing on it, communicator/ros_inference.py:139).
        '''
        h, w = cv_image.shape[:2]
        return self.image_adjust_batch(cv_image[None], (h, w), mode="stretch")

    # ---- added: fused resize (replaces cv2.resize + image_adjust, ros_inference.py:140-141) --------
    def resize_adjust(self, cv_image, size_hw, mode="stretch", swap=False, pad_value=114):
        """One frame at camera resolution -> (1,3,H,W) float32 host array, plus the inverse mapping
        [gain_x, gain_y, pad_left, pad_top] for scaling boxes back (ros_inference.py:100-115)."""
        dev, aff = self.image_adjust_device(cv_image[None] if cv_image.ndim == 3 else cv_image, size_hw, mode, swap, pad_value)
        return self.stager.to_host(dev, "pre_out"), aff

    def image_adjust_batch(self, frames, size_hw=None, mode="stretch", swap=False, pad_value=114, copy=True):
        """frames: (n,h,w,3) uint8 host array/tensor -> (n,3,H,W) float32 host ndarray.
        copy=False hands back a view of the pinned staging buffer (valid until the next call)."""
        if size_hw is None:
            size_hw = tuple(frames.shape[1:3])
        dev, _ = self.image_adjust_device(frames, size_hw, mode, swap, pad_value)
        return self.stager.to_host(dev, "pre_out", copy=copy)

    def image_adjust_device(self, frames, size_hw, mode="stretch", swap=False, pad_value=114, out=None):
        """Same, result left on the device (torch CUDA tensor) for callers that keep going on the GPU."""
        if isinstance(frames, np.ndarray) and frames.dtype != np.uint8:
            raise TypeError("image_adjust: expected a uint8 HxWx3 image")
        shape = tuple(frames.shape)
        src = self.stager.to_device(frames, np.uint8, shape, "pre_in")
        return self.engine.preprocess_images(src, int(size_hw[0]), int(size_hw[1]), mode=mode, swap=swap,
                                             div=self.DIV, pad_value=pad_value, out=out)


class Yolov5preprocess(_ImagePreprocess):
    DIV = 255.0
