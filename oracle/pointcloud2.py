"""CPU restatement of the ROS message handling around the 3D path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

read_points: `np.array(list(point_cloud2.read_points(msg, field_names=("x","y","z","intensity"), skip_nans=True)))`
(communicator/ros_inference3d.py:125). `sensor_msgs.point_cloud2` belongs to ROS (ros/common_msgs, sensor_msgs
1.13.x as shipped with ROS Noetic, the distribution the reference's docker builds on) and is NOT in the reference
tree, nor installable here: **parity unpinned**. Restated from its published source: `_get_struct_fmt` walks the
fields sorted by offset, keeps those named in field_names, pads the gaps; `read_points` unpacks every point with that
struct format in storage order and, with skip_nans, drops a point when any unpacked value is NaN. The tuple order is
therefore the byte-offset order of the fields, not the order of field_names.

conditioning: ros_inference3d.py:127-128 (intensity / max, z + offset) in float64 as the reference runs it.

jsk_rows: the BoundingBoxArray fields of ros_inference3d.py:150-177; the quaternion is pyquaternion 0.9.x's
`Quaternion(axis=[0,0,1], radians=yaw)` (`_from_axis_angle`: unit axis, r = cos(angle/2), i = axis*sin(angle/2));
pyquaternion is not installed here either, the four-line formula is restated.
"""
import math
import struct
from types import SimpleNamespace

import numpy as np

_DATATYPES = {1: ("b", 1), 2: ("B", 1), 3: ("h", 2), 4: ("H", 2), 5: ("i", 4), 6: ("I", 4), 7: ("f", 4), 8: ("d", 8)}


def make_field(name, offset, datatype, count=1):
    return SimpleNamespace(name=name, offset=offset, datatype=datatype, count=count)


def make_cloud(points_by_field, fields, point_step, height=1, row_pad=0, is_bigendian=False, fill=0xAB):
    """Build a PointCloud2-like message from per-field arrays of n = height*width values."""
    n = len(next(iter(points_by_field.values())))
    assert n % height == 0
    width = n // height
    row_step = width * point_step + row_pad
    data = bytearray([fill]) * (row_step * height)
    e = ">" if is_bigendian else "<"
    for f in fields:
        fmt, _ = _DATATYPES[f.datatype]
        vals = points_by_field[f.name]
        for i in range(n):
            v, u = divmod(i, width)
            struct.pack_into(e + fmt, data, v * row_step + u * point_step + f.offset, vals[i])
    return SimpleNamespace(width=width, height=height, point_step=point_step, row_step=row_step, data=bytes(data),
                           is_bigendian=is_bigendian, fields=list(fields))


def selected_fields(fields, field_names=None):
    """The fields read_points unpacks, in the order their values appear in each tuple."""
    return [f for f in sorted(fields, key=lambda f: f.offset) if field_names is None or f.name in field_names]


def _struct_fmt(is_bigendian, fields, field_names=None):
    fmt = ">" if is_bigendian else "<"
    offset = 0
    for f in selected_fields(fields, field_names):
        if offset < f.offset:
            fmt += "x" * (f.offset - offset)
            offset = f.offset
        t, size = _DATATYPES[f.datatype]
        fmt += t * f.count
        offset += f.count * size
    return fmt


def read_points(cloud, field_names=None, skip_nans=False):
    """Generator of tuples, as sensor_msgs.point_cloud2.read_points (uvs=[])."""
    unpack_from = struct.Struct(_struct_fmt(cloud.is_bigendian, cloud.fields, field_names)).unpack_from
    for v in range(cloud.height):
        offset = cloud.row_step * v
        for _ in range(cloud.width):
            p = unpack_from(cloud.data, offset)
            if not (skip_nans and any(math.isnan(pv) for pv in p)):
                yield p
            offset += cloud.point_step


def condition(pc, offset=1.5):
    """ros_inference3d.py:127-128 on the float64 (N,4) array."""
    pc = np.array(pc, dtype=np.float64)
    pc[:, 3] = pc[:, 3] / np.max(pc[:, 3])
    pc[:, 2] = pc[:, 2] + offset
    return pc


def jsk_rows(box_array, scores, labels, label_sel=2, score_thresh=0.5, offset=1.5):
    """(indices, rows): per published box [px, py, pz, ox, oy, oz, ow, dx, dy, dz, value, label] (float64)."""
    indices = np.where((labels == label_sel) & (scores > score_thresh))[0].tolist()
    rows = []
    axis = np.array([0.0, 0.0, 1.0])
    for i in indices:
        theta = float(box_array[i][6]) / 2.0
        q = [math.cos(theta)] + list(axis * math.sin(theta))          # pyquaternion: [w, x, y, z]
        rows.append([float(box_array[i][0]), float(box_array[i][1]), float(box_array[i][2]) - offset,
                     q[1], q[2], q[3], q[0],
                     float(box_array[i][4]), float(box_array[i][3]), float(box_array[i][5]),
                     float(scores[i]), float(int(labels[i]))])
    return indices, np.array(rows, dtype=np.float64).reshape(-1, 12)
