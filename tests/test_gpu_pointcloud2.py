"""GPU parity for the ROS message decode/encode kernels through the C ABI: PointCloud2 unpack
(bit-exact rows, order, count) alone and chained into the voxeliser without a host sync, BoundingBoxArray fields."""
import numpy as np
import pytest
import torch

from oracle import pointcloud2 as opc2
from oracle import voxelize as ov
from triton_client_b200 import synth
from triton_client_b200.clients.preprocess import pointcloud2 as pc2

pytestmark = pytest.mark.gpu

KITTI = ([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4])
XYZI = ("x", "y", "z", "intensity")


def _msg(pts, height=1, row_pad=0, big=False, nan_frac=0.03, seed=0, point_step=32, inten_type=7):
    rng = np.random.default_rng(seed)
    pts = pts.copy()
    n = len(pts)
    pts[rng.random(n) < nan_frac, 0] = np.nan
    pts[rng.random(n) < nan_frac / 2, 3 if inten_type == 7 else 2] = np.nan
    inten = pts[:, 3] if inten_type == 7 else np.clip(pts[:, 3] * 255, 0, 255).astype(np.uint8)
    fields = [opc2.make_field("x", 0, 7), opc2.make_field("y", 4, 7), opc2.make_field("z", 8, 7),
              opc2.make_field("intensity", 16, inten_type), opc2.make_field("ring", 20, 4), opc2.make_field("t", 24, 3)]
    cols = dict(x=pts[:, 0], y=pts[:, 1], z=pts[:, 2], intensity=inten, ring=rng.integers(0, 64, n).astype(np.uint16),
                t=rng.integers(-3000, 3000, n).astype(np.int16))
    return opc2.make_cloud(cols, fields, point_step, height=height, row_pad=row_pad, is_bigendian=big)


@pytest.mark.parametrize("height,row_pad,big,inten_type,names,skip", [
    (1, 0, False, 7, XYZI, True), (16, 10, False, 7, XYZI, True), (4, 3, True, 7, XYZI, True), (1, 0, False, 2, XYZI, True),
    (2, 0, False, 7, None, True), (1, 0, False, 7, ("x", "y", "z"), False), (8, 0, True, 7, ("ring", "t", "y"), False)])
def test_unpack_matches_read_points(engine, height, row_pad, big, inten_type, names, skip):
    pts = synth.lidar_sweep(5, beams=16, azimuths=250)[:3200]
    msg = _msg(pts, height, row_pad, big, seed=height, inten_type=inten_type)
    want = np.array(list(opc2.read_points(msg, field_names=names, skip_nans=skip)), dtype=np.float64)
    got, count = pc2.read_points_device(msg, names, skip, engine=engine)
    k = int(count.item())
    assert k == len(want) and got.shape[0] == 3200
    g = got.cpu().numpy()
    assert np.array_equal(g[:k].astype(np.float64), want, equal_nan=True)
    if k < 3200:
        assert np.isnan(g[k:, :min(3, g.shape[1])]).all() and (g.shape[1] <= 3 or np.isneginf(g[k:, 3:]).all())
    assert np.array_equal(pc2.read_points(msg, names, skip, engine=engine), want, equal_nan=True)


def test_empty_and_all_nan_clouds(engine):
    msg = _msg(np.zeros((0, 4), np.float32))
    got, count = pc2.read_points_device(msg, XYZI, True, engine=engine)
    assert got.shape == (0, 4) and int(count.item()) == 0
    pts = np.full((100, 4), np.nan, np.float32)
    got, count = pc2.read_points_device(_msg(pts, nan_frac=0), XYZI, True, engine=engine)
    assert int(count.item()) == 0 and np.isnan(got.cpu().numpy()[:, :3]).all()


def test_unsupported_layouts_are_errors(engine):
    from triton_client_b200._lib import TcbError
    msg = _msg(synth.lidar_sweep(5, beams=4, azimuths=50))
    msg.fields[0].datatype = 8   # FLOAT64 x
    with pytest.raises(ValueError):
        pc2.read_points_device(msg, XYZI, True, engine=engine)
    with pytest.raises(TcbError):
        engine.pointcloud2_unpack(torch.zeros(3200, dtype=torch.uint8, device="cuda"), 100, 1, 32, 3200, [4, 0], [7, 7])
    with pytest.raises(ValueError):
        engine.pointcloud2_unpack(torch.zeros(100, dtype=torch.uint8, device="cuda"), 100, 1, 32, 3200, [0], [7])


def test_c4_sweep_message_to_voxels_without_host_sync(engine):
    """C4 (64 beams x 1875 azimuths): message bytes -> rows -> conditioning + voxelisation, all on the device; equal to the
    reference pipeline read_points -> intensity/max, z+1.5 (float64) -> filter_pc restated on the CPU."""
    pts = synth.lidar_sweep(11)
    pts = pts[:len(pts) // 64 * 64]
    pts[:, 3] *= 100.0
    msg = _msg(pts, height=64, nan_frac=0.02, seed=3)
    host = np.array(list(opc2.read_points(msg, field_names=XYZI, skip_nans=True)))
    d = ov.filter_pc_pointpillar(ov.condition_points(host.astype(np.float32), 1.5), *KITTI, 32, 40000)
    dev, count = pc2.read_points_device(msg, XYZI, True, engine=engine)
    assert dev.shape[0] == len(pts) and int(count.item()) == len(host)
    v, c, n, m = engine.voxelize(dev, KITTI[0], KITTI[1], 32, 40000, range_mask=True, z_offset=1.5, intensity_col=3, nfeat=4)
    m = int(m.item())
    assert m == len(d["voxel_coords"]) and m > 5000
    assert np.array_equal(c[:m].cpu().numpy(), d["voxel_coords"]) and np.array_equal(n[:m].cpu().numpy(), d["voxel_num_points"])
    assert np.array_equal(v[:m].cpu().numpy(), d["voxels"])


def test_bounding_box_array_fields(engine):
    rng = np.random.default_rng(2)
    boxes = synth.random_boxes7(400, 9, extent=40)
    boxes[:, 6] = rng.uniform(-7, 7, 400)
    scores = rng.uniform(0, 1, 400).astype(np.float32)
    scores[::9] = 0.5                                    # the cut is strict
    labels = rng.integers(1, 4, 400)
    idx, want = opc2.jsk_rows(boxes, scores, labels)
    rows, gi, cnt = pc2.jsk_box_fields(boxes, scores, labels, engine=engine)
    k = int(cnt.item())
    assert k == len(idx) > 20 and gi[:k].cpu().tolist() == idx
    got = rows[:k].cpu().numpy()
    assert np.array_equal(got[:, [0, 1, 2, 3, 4, 7, 8, 9, 10, 11]], want[:, [0, 1, 2, 3, 4, 7, 8, 9, 10, 11]])
    np.testing.assert_allclose(got[:, 5:7], want[:, 5:7], rtol=0, atol=2.5e-16)     # sin/cos: libm vs CUDA, one ulp
    rows, gi, cnt = pc2.jsk_box_fields(boxes[:0], scores[:0], labels[:0], engine=engine)
    assert int(cnt.item()) == 0
