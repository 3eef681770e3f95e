"""Oracle checks for voxelisation / decoration: C restatement vs its pure-Python twin, driver semantics,
and the decoration restatement vs golden vectors produced by the reference's own PillarVFE code."""
import numpy as np
import pytest

from conftest import golden
from oracle import ref_harness
from oracle import voxelize as ov
from triton_client_b200 import synth

KITTI = ([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4])


@pytest.mark.parametrize("overflow", ["continue", "break"])
def test_c_voxeliser_equals_python_twin(overflow):
    pts = synth.lidar_sweep(41, beams=16, azimuths=300)
    for max_vox in (50, 100000):
        a = ov.points_to_voxel(pts, *KITTI, 8, max_vox, overflow)
        b = ov.points_to_voxel_py(pts, *KITTI, 8, max_vox, overflow)
        assert all(np.array_equal(x, y) for x, y in zip(a, b))


def test_first_seen_order_and_padding():
    pts = np.array([[1.0, 1.0, 0, 9], [5.0, 5.0, 0, 8], [1.01, 1.01, 0, 7], [100, 0, 0, 6], [1.02, 1.0, 0, 5]], np.float32)
    v, c, n = ov.points_to_voxel(pts, [0, 0, -1, 10, 10, 1], [1, 1, 2], 2, 10)
    assert c.tolist() == [[0, 1, 1], [0, 5, 5]] and n.tolist() == [2, 1]
    assert v[0, :, 3].tolist() == [9, 7] and v[1, 1].tolist() == [0, 0, 0, 0]


def test_driver_semantics():
    pts = np.array([[0.0, -39.68, 5.0, 1], [69.12, 39.68, 0, 1], [69.13, 0, 0, 1], [10, 10, 0.5, 1]], np.float32)
    d = ov.filter_pc_pointpillar(pts, *KITTI, 32, 100)
    assert len(d["points"]) == 3                      # inclusive x/y mask, z not tested (common_utils.py:60-63)
    assert d["voxel_coords"].shape[1] == 4 and (d["voxel_coords"][:, 0] == 0).all()
    assert d["voxels"].shape[0] == 1                  # z=5 and the x==69.12 edge are rejected by the grid test
    r = ov.filter_pc_det3d(pts, [-51.2, -51.2, -5, 51.2, 51.2, 3], [0.2, 0.2, 8], 20, 100)
    assert r[0].shape[2] == 5 and r[0].dtype == np.float32 and r[1].dtype == np.int32


def test_golden_pillar_decoration():
    g = golden("pillar_vfe")
    for k in range(2):
        rng, vs, p = g[f"range{k}"].tolist(), g[f"vsize{k}"].tolist(), int(g[f"maxpts{k}"])
        voxels, coords, num = ov.points_to_voxel(g[f"points{k}"], rng, vs, p, 400)
        feats = ov.pillar_decorate(voxels, np.pad(coords, ((0, 0), (1, 0))), num, vs, np.array(rng, np.float32) if k == 0 else rng)
        assert np.array_equal(feats, g[f"feats{k}"])


@pytest.mark.skipif(not ref_harness.available(), reason="reference not mounted (GPU box)")
def test_mask_against_live_reference():
    cu = ref_harness.common_utils()
    pts = synth.lidar_sweep(42, beams=8, azimuths=100)
    rng = np.array(KITTI[0], np.float32)
    assert np.array_equal(cu.mask_points_by_range(pts, rng), ov.mask_points_by_range(pts, rng))


def test_golden_dynamic_pillar_vfe():
    """DynamicPillarVFE.forward run unmodified (oracle/make_golden.py: golden_dynamic_vfe) vs the restatement."""
    g = golden("dynamic_vfe")
    f, inv, c = ov.dynamic_pillar_vfe(g["points"], g["vsize"].tolist(), g["range"].tolist(), g["grid"].tolist())
    assert np.array_equal(inv, g["unq_inv"]) and np.array_equal(c, g["voxel_coords"])
    assert np.array_equal(f, g["feats"])
