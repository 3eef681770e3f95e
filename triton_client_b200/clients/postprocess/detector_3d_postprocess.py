"""Drop-in for clients/postprocess/detector_3d_postprocess.py (PointPillarPostprocess). extract_boxes keeps the
reference's contract (three wire tensors -> dict of object arrays). The decode + rotated NMS that the reference
leaves to the Triton Python backend (examples/pointpillar_kitti/1/model.py:163 -> pcdet) are ADDED methods that run
on the GPU: decode_pointpillars, decode_centerpoint, rotated_nms."""
import numpy as np
import torch

from .base_postprocess import Postprocess

# (class, threshold) in the order remove_low_score_nu concatenates them (detector_3d_postprocess.py:103-127);
# motorcycle (6, 0.15) indices are computed by the reference but never appended.
NUSC_GROUPS = [(0, 0.4), (8, 0.1), (7, 0.15), (3, 0.3), (2, 0.4), (9, 0.1), (4, 0.4), (5, 0.4), (1, 0.4)]

# data/pointpillar.yaml:74-147 of the reference (anchor head + post-processing of the served KITTI model)
KITTI_PP = dict(anchor_range=[0, -39.68, -3, 69.12, 39.68, 1], anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]],
                anchor_bottoms=[-1.78, -0.6, -0.6], anchor_rotations=[[0, 1.57]] * 3, dir_offset=0.78539, dir_limit_offset=0.0,
                num_dir_bins=2, score_thresh=0.1, nms_thresh=0.01, pre_max=4096, post_max=500)
# OPCD/tools/cfgs/nuscenes_models/cbgs_dyn_pp_centerpoint.yaml:47-90
NUSC_CP = dict(class_id_map=[[0], [1, 2], [3, 4], [5], [6, 7], [8, 9]], feature_map_stride=4, voxel_size=[0.2, 0.2, 8.0],
               point_range=[-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], limit_range=[-61.2, -61.2, -10.0, 61.2, 61.2, 10.0], k=500,
               score_thresh=0.1, nms_thresh=0.2, pre_max=1000, post_max=83)


class PointPillarPostprocess(Postprocess):
    def __init__(self, engine=None):
        super().__init__(engine)

    def postprocess(self):
        pass

    def load_class_names(self, namesfile='./data/nuScenes.names'):
        with open(namesfile, 'r') as fp:
            return [line.rstrip() for line in fp.readlines()]

    # ---- reference surface -----------------------------------------------------------------------
    def extract_boxes(self, prediction):
        self.boxes = np.reshape(self.deserialize_bytes_float(prediction.raw_output_contents[0]), prediction.outputs[0].shape)
        self.scores = np.reshape(self.deserialize_bytes_float(prediction.raw_output_contents[1]), prediction.outputs[1].shape)
        self.labels = np.reshape(self.deserialize_bytes_int(prediction.raw_output_contents[2]), prediction.outputs[2].shape)
        self.output = {'pred_boxes': self.boxes, 'pred_scores': self.scores, 'pred_labels': self.labels}
        return self.output

    def get_annotations_indices(self, types, thresh, label_preds, scores):
        idx, _ = self._class_filter(label_preds, scores, [(types, thresh)])
        return idx

    def remove_low_score_nu(self, predictions, thresh=None):
        idx, _ = self._class_filter(predictions["labels"], predictions["scores"], NUSC_GROUPS)
        return {k: v[idx] for k, v in predictions.items() if k != 'metadata'}

    def _class_filter(self, labels, scores, groups):
        lab = torch.as_tensor(np.asarray(labels).astype(np.int64)).to(self.engine.device)
        sc = torch.as_tensor(np.asarray(scores).astype(np.float32)).to(self.engine.device)
        out, cnt = self.engine.class_score_filter(lab, sc, groups)
        n = int(cnt.item())
        return out[:n].cpu().tolist(), n

    # ---- added: the server-side decode, client-side on the GPU ---------------------------------------
    def decode_pointpillars(self, cls_preds, box_preds, dir_cls_preds=None, **cfg):
        """One frame of AnchorHeadSingle outputs (H,W,A*nc / A*7 / A*2; numpy or tensors) -> the reference's
        {'pred_boxes','pred_scores','pred_labels'} (CPU tensors), via anchors + ResidualCoder + direction bins + NMS."""
        c = dict(KITTI_PP)
        c.update(cfg)
        dev = self.engine.device
        t = lambda x: None if x is None else torch.as_tensor(x, dtype=torch.float32).to(dev)
        ob, osc, ol, cnt = self.engine.pp_anchor_post(t(cls_preds), t(box_preds), t(dir_cls_preds), **c)
        n = int(cnt.item())
        return {'pred_boxes': ob[:n].cpu(), 'pred_scores': osc[:n].cpu(), 'pred_labels': ol[:n].cpu()}

    def decode_centerpoint(self, heads, **cfg):
        """heads: list of dicts (hm, center, center_z, dim, rot[, vel]) of one frame -> the reference's dict."""
        c = dict(NUSC_CP)
        c.update(cfg)
        dev = self.engine.device
        hd = [{k: torch.as_tensor(v, dtype=torch.float32).to(dev) for k, v in h.items()} for h in heads]
        cmap = c.pop("class_id_map")
        ob, osc, ol, cnt = self.engine.centerpoint_post(hd, cmap[:len(hd)], c.pop("feature_map_stride"), c.pop("voxel_size"),
                                                        c.pop("point_range"), c.pop("limit_range"), **c)
        n = int(cnt.item())
        return {'pred_boxes': ob[:n].cpu(), 'pred_scores': osc[:n].cpu(), 'pred_labels': ol[:n].cpu()}

    def multi_classes_nms(self, cls_scores, box_preds, nms_thresh, pre_max=4096, post_max=500, score_thresh=None, nms_type="nms_gpu"):
        """model_nms_utils.multi_classes_nms (the MULTI_CLASSES_NMS branch of post_processing) -> CPU tensors."""
        dev = self.engine.device
        ps, pl, pb = self.engine.multi_classes_nms(torch.as_tensor(cls_scores, dtype=torch.float32).to(dev), torch.as_tensor(box_preds, dtype=torch.float32).to(dev),
                                                   nms_thresh, pre_max, post_max, score_thresh, nms_type)
        return ps.cpu(), pl.cpu(), pb.cpu()

    def rotated_nms(self, boxes, scores, thresh, pre_max=4096, post_max=500, nms_type="nms_gpu"):
        dev = self.engine.device
        keep, cnt = self.engine.rotated_nms(torch.as_tensor(boxes, dtype=torch.float32).to(dev), torch.as_tensor(scores, dtype=torch.float32).to(dev),
                                            thresh, pre_max, post_max, nms_type)
        return keep[:int(cnt.item())].cpu().long()
