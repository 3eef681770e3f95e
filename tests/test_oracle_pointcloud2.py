"""Host-side checks of the PointCloud2 / BoundingBoxArray restatement (parity unpinned: the ROS
library it restates is not in the reference tree). The restatement is checked against numpy's own view of the bytes."""
import numpy as np

from oracle import pointcloud2 as pc2


def _velodyne_like(n, seed, height=1, row_pad=0, big=False):
    rng = np.random.default_rng(seed)
    x, y, z = (rng.normal(0, 20, n).astype(np.float32) for _ in range(3))
    inten = rng.uniform(0, 255, n).astype(np.float32)
    ring = rng.integers(0, 64, n).astype(np.uint16)
    x[rng.random(n) < 0.05] = np.nan
    inten[rng.random(n) < 0.02] = np.nan
    fields = [pc2.make_field("x", 0, 7), pc2.make_field("y", 4, 7), pc2.make_field("z", 8, 7), pc2.make_field("intensity", 16, 7),
              pc2.make_field("ring", 20, 4)]
    msg = pc2.make_cloud(dict(x=x, y=y, z=z, intensity=inten, ring=ring), fields, 22, height=height, row_pad=row_pad, is_bigendian=big)
    return msg, np.stack([x, y, z, inten], 1), ring


def test_read_points_matches_a_numpy_view_of_the_bytes():
    for (height, pad, big) in [(1, 0, False), (4, 6, False), (2, 3, True)]:
        msg, xyzi, ring = _velodyne_like(400, 7 + height, height, pad, big)
        got = np.array(list(pc2.read_points(msg, field_names=("x", "y", "z", "intensity"), skip_nans=True)))
        keep = ~np.isnan(xyzi).any(1)
        assert got.dtype == np.float64 and np.array_equal(got, xyzi[keep].astype(np.float64))
        allf = np.array(list(pc2.read_points(msg, skip_nans=False)))
        assert allf.shape == (400, 5) and np.array_equal(allf[:, 4], ring.astype(np.float64))
        # tuple order is offset order, not the order of field_names
        swapped = np.array(list(pc2.read_points(msg, field_names=("intensity", "x"), skip_nans=False)))
        assert np.array_equal(np.nan_to_num(swapped), np.nan_to_num(xyzi[:, [0, 3]].astype(np.float64)))


def test_condition_and_jsk_rows():
    msg, xyzi, _ = _velodyne_like(300, 3)
    pts = np.array(list(pc2.read_points(msg, field_names=("x", "y", "z", "intensity"), skip_nans=True)))
    c = pc2.condition(pts)
    assert np.isclose(c[:, 3].max(), 1.0) and np.allclose(c[:, 2], pts[:, 2] + 1.5)
    rng = np.random.default_rng(1)
    boxes = rng.normal(0, 10, (50, 7)).astype(np.float32)
    scores = rng.uniform(0, 1, 50).astype(np.float32)
    labels = rng.integers(1, 4, 50)
    idx, rows = pc2.jsk_rows(boxes, scores, labels)
    assert idx == [i for i in range(50) if labels[i] == 2 and scores[i] > 0.5] and rows.shape == (len(idx), 12)
    for k, i in enumerate(idx):
        assert np.isclose(rows[k, 3:7] @ rows[k, 3:7], 1.0) and rows[k, 3] == 0 and rows[k, 4] == 0
        half = float(boxes[i, 6]) / 2.0
        assert rows[k, 5] == np.sin(half) and rows[k, 6] == np.cos(half)
        assert rows[k, 7] == boxes[i, 4] and rows[k, 8] == boxes[i, 3] and rows[k, 2] == float(boxes[i, 2]) - 1.5
