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

    def image_adjust_batch(self, frames, size_hw=None, mode="stretch", swap=False, pad_value=114, copy=True, chunks=16):
        """frames: (n,h,w,3) uint8 host array/tensor -> (n,3,H,W) float32 host ndarray.
        copy=False hands back a view of the pinned staging buffer (valid until the next call).
        Large batches are cut into `chunks` pieces that flow through three streams (H2D, kernel, D2H), so the
        upload of chunk k+1 overlaps the download of chunk k on the full-duplex PCIe link."""
        if size_hw is None:
            size_hw = tuple(frames.shape[1:3])
        n = int(frames.shape[0])
        chunks = min(int(chunks), n // 2)      # at least two frames per piece (measured on 64 frames: 8 pieces 6.53 ms, 16: 6.35, 32: 6.34)
        if chunks < 2:
            dev, _ = self.image_adjust_device(frames, size_hw, mode, swap, pad_value)
            return self.stager.to_host(dev, "pre_out", copy=copy)
        return self._pipelined_batch(frames, size_hw, mode, swap, pad_value, copy, chunks)

    def image_adjust_batch_async(self, frames, size_hw=None, mode="stretch", swap=False, pad_value=114, copy=True, chunks=16):
        """image_adjust_batch for a pipelined caller: every copy and kernel of the batch is queued and the call returns a
        handle at once; `.result()` waits and returns the (n,3,H,W) float32 host array. Work queued afterwards on other
        streams (an extract_boxes_async, the next batch's image_adjust_batch_async) shares the GPU and the PCIe link with
        it. At most two batches in flight: results alternate between two host buffers."""
        if size_hw is None:
            size_hw = tuple(frames.shape[1:3])
        n = int(frames.shape[0])
        chunks = min(int(chunks), n // 2)
        if chunks < 2:
            return _Ready(self.image_adjust_batch(frames, size_hw, mode, swap, pad_value, copy, chunks))
        return _Deferred(self._pipelined_batch(frames, size_hw, mode, swap, pad_value, copy, chunks, defer=True))

    def _pipelined_batch(self, frames, size_hw, mode, swap, pad_value, copy, chunks, defer=False):
        eng, st = self.engine, self.stager
        dev = eng.device
        n, h, w, _ = (int(v) for v in frames.shape)
        oh, ow = int(size_hw[0]), int(size_hw[1])
        if isinstance(frames, np.ndarray):
            src = torch.from_numpy(np.ascontiguousarray(frames))
        else:
            src = frames.contiguous()
        # deferred calls alternate between two sets of host buffers, so that the next batch can be queued while the
        # previous one's uploads are still reading their staging buffer and the caller still reads the previous result
        # (copy=False views stay valid until the next call but one)
        self._out_turn = (getattr(self, "_out_turn", 0) + 1) & 1 if defer else 0
        staged = not src.is_pinned()
        if staged:   # pageable source: stage it once (a pinned caller buffer avoids this memcpy)
            if getattr(self, "_in_done", None) is None:
                self._in_done = [None, None]
            if self._in_done[self._out_turn] is not None:
                self._in_done[self._out_turn].synchronize()    # the uploads that last read this staging buffer have ended
            pin = st.pinned("pre_in_host" if self._out_turn == 0 else "pre_in_host_b", src.numel()).view(n, h, w, 3)
            pin.copy_(src)
            src = pin
        host_out = st.pinned("pre_pipe_a" if self._out_turn == 0 else "pre_pipe_b", n * 3 * oh * ow * 4).view(torch.float32).view(n, 3, oh, ow)
        # The padding of a letterboxed / padded model input is one constant per geometry: once a result buffer holds it (a
        # first, full download), later batches of the same geometry only bring back the rows and columns that carry the
        # frame (configs[1]: 360 of 640 rows per plane) — the host array ends up with the same bytes as a full download.
        rect = eng.content_rect(mode, h, w, oh, ow)
        geo = (host_out.data_ptr(), n, h, w, oh, ow, mode, int(pad_value), bool(swap), float(self.DIV))
        if getattr(self, "_filled", None) is None:
            self._filled = [None, None]
        partial = rect != (0, 0, ow, oh) and self._filled[self._out_turn] == geo
        self._filled[self._out_turn] = None          # until this call's downloads have been queued in full
        if getattr(self, "_streams", None) is None:
            self._streams = (torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev))
        s_in, s_cmp, s_out = self._streams
        per = (n + chunks - 1) // chunks
        key = (per, h, w, oh, ow)
        if getattr(self, "_pipe_key", None) != key:
            self._pipe_in = [torch.empty((per, h, w, 3), dtype=torch.uint8, device=dev) for _ in range(2)]
            self._pipe_out = [torch.empty((per, 3, oh, ow), dtype=torch.float32, device=dev) for _ in range(2)]
            self._pipe_key = key
        cur = torch.cuda.current_stream(dev)
        for s in self._streams:
            s.wait_stream(cur)
        # a previous deferred call may still be running: its kernels read the input buffers, its downloads the output buffers
        s_in.wait_stream(s_cmp)
        s_cmp.wait_stream(s_out)
        ev_cmp, ev_out = [], []
        for k in range(chunks):
            a, b = k * per, min(n, (k + 1) * per)
            if a >= b:
                break
            m = b - a
            din, dout = self._pipe_in[k % 2], self._pipe_out[k % 2]
            with torch.cuda.stream(s_in):
                if k >= 2:
                    s_in.wait_event(ev_cmp[k - 2])      # the kernel that last read this input buffer is done
                din[:m].copy_(src[a:b], non_blocking=True)
                e_in = torch.cuda.Event()
                e_in.record(s_in)
            with torch.cuda.stream(s_cmp):
                s_cmp.wait_event(e_in)
                if k >= 2:
                    s_cmp.wait_event(ev_out[k - 2])     # the download that last read this output buffer is done
                eng.preprocess_images(din[:m], oh, ow, mode=mode, swap=swap, div=self.DIV, pad_value=pad_value, out=dout[:m])
                e_c = torch.cuda.Event()
                e_c.record(s_cmp)
                ev_cmp.append(e_c)
            with torch.cuda.stream(s_out):
                s_out.wait_event(e_c)
                if partial:
                    eng.copy_rect_d2h(dout[:m], host_out[a:b], rect)
                else:
                    host_out[a:b].copy_(dout[:m], non_blocking=True)
                e_o = torch.cuda.Event()
                e_o.record(s_out)
                ev_out.append(e_o)
        done = ev_out[-1]
        self._filled[self._out_turn] = geo
        self.last_d2h_bytes = n * 3 * (rect[2] * rect[3] if partial else oh * ow) * 4
        if staged:
            self._in_done[self._out_turn] = e_in

        def finish():
            done.synchronize()             # this call's last download (a later call's work may already be queued behind it)
            cur.wait_stream(s_cmp)
            res = host_out.numpy()
            if copy:
                return res.copy()
            res.flags.writeable = False      # a view of the staging buffer: its padding is reused by the next batches
            return res
        return finish if defer else finish()

    def image_adjust_device(self, frames, size_hw, mode="stretch", swap=False, pad_value=114, out=None):
        """Same, result left on the device (torch CUDA tensor) for callers that keep going on the GPU."""
        if isinstance(frames, np.ndarray) and frames.dtype != np.uint8:
            raise TypeError("image_adjust: expected a uint8 HxWx3 image")
        shape = tuple(frames.shape)
        src = self.stager.to_device(frames, np.uint8, shape, "pre_in")
        return self.engine.preprocess_images(src, int(size_hw[0]), int(size_hw[1]), mode=mode, swap=swap,
                                             div=self.DIV, pad_value=pad_value, out=out)



class _Ready:
    def __init__(self, value):
        self._value = value

    def result(self):
        return self._value


class _Deferred:
    def __init__(self, finish):
        self._finish, self._value = finish, None

    def result(self):
        if self._finish is not None:
            self._value, self._finish = self._finish(), None
        return self._value


class Yolov5preprocess(_ImagePreprocess):
    DIV = 255.0
