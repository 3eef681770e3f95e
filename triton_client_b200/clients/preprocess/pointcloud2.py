"""PointCloud2 message -> point rows on the GPU, and detections -> BoundingBoxArray fields.

The reference unpacks every LiDAR message with a Python generator, one struct.unpack per point
(communicator/ros_inference3d.py:125: `np.array(list(point_cloud2.read_points(msg, field_names=(...), skip_nans=True)))`),
conditions the array (:127-128) and hands it to `filter_pc`; after the model call it fills one BoundingBox message per
kept detection (:150-177). Here the message bytes go to the device once and stay there:

    pts, count = read_points_device(msg, ("x", "y", "z", "intensity"), skip_nans=True)      # CUDA [width*height, 4] f32
    voxels, coords, num, m = preprocess.filter_pc_device(pts, z_offset=1.5, intensity_col=3) # tcb_voxelize, no host sync
    rows, idx, n = jsk_box_fields(pred_boxes, pred_scores, pred_labels)                      # CUDA [n,12] f64

`msg` is anything with the sensor_msgs/PointCloud2 attributes (width, height, point_step, row_step, data,
is_bigendian, fields[i].name/.offset/.datatype/.count). No CPU path: without the CUDA library these raise.
"""
import numpy as np
import torch

from .. import wire

EXACT_IN_F32 = {1, 2, 3, 4, 7}   # INT8, UINT8, INT16, UINT16, FLOAT32


def _selected(fields, field_names):
    sel = [f for f in sorted(fields, key=lambda f: f.offset) if field_names is None or f.name in field_names]
    for f in sel:
        if getattr(f, "count", 1) != 1:
            raise ValueError(f"read_points: field {f.name!r} has count {f.count}; only scalar fields are supported")
        if f.datatype not in EXACT_IN_F32:
            raise ValueError(f"read_points: field {f.name!r} has datatype {f.datatype}, which float32 rows cannot hold exactly")
    return sel


def read_points_device(msg, field_names=None, skip_nans=False, engine=None):
    """Returns (points, count): CUDA float32 [width*height, k] whose first `count` rows are the points read_points
    yields (columns in ascending field offset, like its tuples) and whose other rows are NaN/-inf fillers that
    tcb_voxelize ignores; count is a CUDA int32 [1]."""
    eng = engine or wire.default_engine()
    sel = _selected(msg.fields, field_names)
    raw = msg.data
    if isinstance(raw, torch.Tensor):
        dev = raw.to(eng.device, non_blocking=True).view(torch.uint8).reshape(-1)
    else:
        host = torch.frombuffer(bytearray(raw) if isinstance(raw, (bytes, memoryview)) else raw, dtype=torch.uint8) if len(raw) else torch.zeros(0, dtype=torch.uint8)
        dev = host.to(eng.device)
    return eng.pointcloud2_unpack(dev, msg.width, msg.height, msg.point_step, msg.row_step, [f.offset for f in sel],
                                  [f.datatype for f in sel], msg.is_bigendian, skip_nans)


def read_points(msg, field_names=None, skip_nans=False, engine=None):
    """Host twin of `np.array(list(point_cloud2.read_points(...)))`: float64 (N, k)."""
    pts, count = read_points_device(msg, field_names, skip_nans, engine)
    n = int(count.item())
    return pts[:n].cpu().numpy().astype(np.float64)


def jsk_box_fields(pred_boxes, pred_scores, pred_labels, label=2, score_thresh=0.5, offset=1.5, engine=None):
    """Fields of the BoundingBoxArray the reference publishes (ros_inference3d.py:150-177). Inputs as extract_boxes
    returns them (numpy or tensors). Returns CUDA (rows f64 [n,12], indices int32 [n], count int32 [1]); a row is
    [position x, y, z; orientation x, y, z, w; dimensions x, y, z; value; label]."""
    eng = engine or wire.default_engine()
    t = lambda a, dt: (a if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a))).to(eng.device).to(dt)  # noqa: E731
    return eng.boxes_to_jsk(t(pred_boxes, torch.float32), t(pred_scores, torch.float32), t(pred_labels, torch.int64), label, score_thresh, offset)
