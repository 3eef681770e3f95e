"""Mirror of tools/yolo_layer.py: yolo_forward_dynamic (:148-288) and YoloLayer (:290-321) on the GPU."""
import torch

from ..clients import wire


def yolo_forward_dynamic(output, conf_thresh, num_classes, anchors, num_anchors, scale_x_y, only_objectness=1, validation=False):
    """output [B, A*(5+nc), H, W]; anchors already divided by the stride. Returns boxes [B,A*H*W,1,4], confs [B,A*H*W,nc]."""
    eng = wire.default_engine()
    t = torch.as_tensor(output, dtype=torch.float32).to(eng.device)
    return eng.yolo_head_decode_v4([t], [list(anchors)[:2 * num_anchors]], num_classes, scale_x_y)


class YoloLayer(torch.nn.Module):
    def __init__(self, anchor_mask=[], num_classes=0, anchors=[], num_anchors=1, stride=32, model_out=False):
        super().__init__()
        self.anchor_mask, self.num_classes, self.anchors, self.num_anchors = anchor_mask, num_classes, anchors, num_anchors
        self.anchor_step = len(anchors) // num_anchors
        self.thresh, self.stride, self.scale_x_y, self.model_out = 0.6, stride, 1, model_out

    def forward(self, output, target=None):
        if self.training:
            return output
        masked = []
        for m in self.anchor_mask:
            masked += self.anchors[m * self.anchor_step:(m + 1) * self.anchor_step]
        masked = [a / self.stride for a in masked]
        return yolo_forward_dynamic(output, self.thresh, self.num_classes, masked, len(self.anchor_mask), scale_x_y=self.scale_x_y)
