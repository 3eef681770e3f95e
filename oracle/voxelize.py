"""Oracle: LiDAR conditioning, range mask, hard voxelisation, pillar decoration.
Test infrastructure only. The voxeliser itself is the C restatement in oracle/c/voxelize.c (PARITY
UNPINNED: spconv / det3d are not vendored); the driver semantics around it are pinned by in-tree code."""
import ctypes
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "c", "liboracle.so")
_lib = None


def clib():
    global _lib
    if _lib is None:
        src = [os.path.join(_HERE, "c", f) for f in os.listdir(os.path.join(_HERE, "c")) if f.endswith(".c")]
        if not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src):
            subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", "-o", _SO] + src + ["-lm"])
        _lib = ctypes.CDLL(_SO)
    return _lib


# --- reference: communicator/ros_inference3d.py:127-128 (bag_inference3d.py:124-126) ----------------
def condition_points(pc, z_offset=1.5):
    """intensity /= max(intensity); z += offset — in float32 (the build's fixed precision)."""
    pc = np.array(pc, dtype=np.float32, copy=True)
    pc[:, 3] = pc[:, 3] / np.max(pc[:, 3])
    pc[:, 2] = pc[:, 2] + np.float32(z_offset)
    return pc


# --- reference: OPCD/pcdet/utils/common_utils.py:60-63 ----------------------------------------------
def mask_points_by_range(points, limit_range):
    return (points[:, 0] >= limit_range[0]) & (points[:, 0] <= limit_range[3]) \
        & (points[:, 1] >= limit_range[1]) & (points[:, 1] <= limit_range[4])


def points_to_voxel(points, point_range, voxel_size, max_pts, max_vox, overflow="continue", f64_math=False):
    """Returns (voxels [M,P,F] f32, coords [M,3] i32 zyx, num_points [M] i32) like VoxelGenerator.generate."""
    pts = np.ascontiguousarray(points, dtype=np.float32)
    n, f = pts.shape
    voxels = np.empty((max_vox, max_pts, f), np.float32)
    coords = np.zeros((max_vox, 3), np.int32)
    num = np.empty((max_vox,), np.int32)
    rng = (ctypes.c_double * 6)(*[float(v) for v in point_range])
    vs = (ctypes.c_double * 3)(*[float(v) for v in voxel_size])
    m = clib().oracle_points_to_voxel(pts.ctypes.data_as(ctypes.c_void_p), n, f, f, rng, vs, int(max_pts), int(max_vox),
                                      1 if overflow == "break" else 0, 1 if f64_math else 0,
                                      voxels.ctypes.data_as(ctypes.c_void_p), coords.ctypes.data_as(ctypes.c_void_p),
                                      num.ctypes.data_as(ctypes.c_void_p))
    assert m >= 0
    return voxels[:m].copy(), coords[:m].copy(), num[:m].copy()


def points_to_voxel_py(points, point_range, voxel_size, max_pts, max_vox, overflow="continue"):
    """Pure-Python twin of the C loop (small inputs only) — cross-checks the C restatement."""
    pts = np.asarray(points, np.float32)
    r = np.asarray(point_range, np.float32)
    vs = np.asarray(voxel_size, np.float32)
    grid = np.round((r[3:] - r[:3]) / vs).astype(np.int64)
    f = pts.shape[1]
    voxels = np.zeros((max_vox, max_pts, f), np.float32)
    coords = np.zeros((max_vox, 3), np.int32)
    num = np.zeros((max_vox,), np.int32)
    lut = {}
    m = 0
    for i in range(pts.shape[0]):
        c = np.floor((pts[i, :3] - r[:3]) / vs)
        if np.any(c < 0) or np.any(c >= grid):
            continue
        key = (int(c[2]), int(c[1]), int(c[0]))
        v = lut.get(key, -1)
        if v == -1:
            if m >= max_vox:
                if overflow == "break":
                    break
                continue
            v = m
            m += 1
            lut[key] = v
            coords[v] = key
        if num[v] < max_pts:
            voxels[v, num[v]] = pts[i]
            num[v] += 1
    return voxels[:m], coords[:m], num[:m]


# --- reference: clients/preprocess/preprocess_3d.py:33-51 (PointpillarPreprocess.filter_pc) ----------
def filter_pc_pointpillar(points, point_range, voxel_size, max_pts, max_vox):
    """PointFeatureEncoder (identity on x,y,z,intensity) -> xy range mask -> spconv voxeliser ('continue')
    -> batch column prepended to the coords (preprocess_3d.py:50)."""
    pts = np.asarray(points, np.float32)
    pts = pts[mask_points_by_range(pts, np.asarray(point_range, np.float32))]
    voxels, coords, num = points_to_voxel(pts, point_range, voxel_size, max_pts, max_vox, "continue")
    coords = np.pad(coords, ((0, 0), (1, 0)), mode='constant', constant_values=0)
    return {"points": pts, "frame_id": 1, "use_lead_xyz": True, "voxels": voxels, "voxel_coords": coords, "voxel_num_points": num}


# --- reference: clients/preprocess/voxelize.py:28-48 (det3DPreprocess.filter_pc) ---------------------
def filter_pc_det3d(points, point_range, voxel_size, max_pts, max_vox, f64_math=False):
    """zero feature column appended (:38-39), det3d voxeliser ('break'), voxels cast to f32 (:44),
    batch column prepended (:46-47). No range mask on this path."""
    pts = np.concatenate((np.asarray(points, np.float32), np.zeros((len(points), 1), np.float32)), axis=1)
    voxels, coords, num = points_to_voxel(pts, point_range, voxel_size, max_pts, max_vox, "break", f64_math)
    coords = np.concatenate((np.zeros((coords.shape[0], 1), np.int32), coords), axis=1)
    return [voxels.astype(np.float32), coords, num]


# --- reference: OPCD/pcdet/models/backbones_3d/vfe/pillar_vfe.py:76-81,86-92,94-118 ------------------
def pillar_decorate(voxels, coords, num_points, voxel_size, point_range):
    """PillarVFE.forward up to (not including) the PFN layers: (M,P,F) -> (M,P,F+6), same torch ops."""
    voxel_features = torch.as_tensor(voxels)
    voxel_num_points = torch.as_tensor(num_points)
    coords = torch.as_tensor(coords)
    voxel_x, voxel_y, voxel_z = voxel_size[0], voxel_size[1], voxel_size[2]
    x_offset = voxel_x / 2 + point_range[0]
    y_offset = voxel_y / 2 + point_range[1]
    z_offset = voxel_z / 2 + point_range[2]
    points_mean = voxel_features[:, :, :3].sum(dim=1, keepdim=True) / voxel_num_points.type_as(voxel_features).view(-1, 1, 1)
    f_cluster = voxel_features[:, :, :3] - points_mean
    f_center = torch.zeros_like(voxel_features[:, :, :3])
    f_center[:, :, 0] = voxel_features[:, :, 0] - (coords[:, 3].to(voxel_features.dtype).unsqueeze(1) * voxel_x + x_offset)
    f_center[:, :, 1] = voxel_features[:, :, 1] - (coords[:, 2].to(voxel_features.dtype).unsqueeze(1) * voxel_y + y_offset)
    f_center[:, :, 2] = voxel_features[:, :, 2] - (coords[:, 1].to(voxel_features.dtype).unsqueeze(1) * voxel_z + z_offset)
    features = torch.cat([voxel_features, f_cluster, f_center], dim=-1)
    voxel_count = features.shape[1]
    actual_num = torch.unsqueeze(voxel_num_points, 1)
    max_num = torch.arange(voxel_count, dtype=torch.int).view([1, -1])
    mask = (actual_num.int() > max_num)
    mask = torch.unsqueeze(mask, -1).type_as(voxel_features)
    features *= mask
    return features.numpy()


# --- reference: OPCD/pcdet/models/backbones_3d/vfe/dynamic_pillar_vfe.py:90-142 (up to the PFN layers) ------
def dynamic_pillar_vfe(points, voxel_size, point_cloud_range, grid_size):
    """points (N, 1+F) rows (batch_idx, x, y, z, ...). torch_scatter (absent here) is restated: scatter_mean =
    index_add of the rows in input order / counts. Returns (features (N', F+6), unq_inv, voxel_coords (U,4))."""
    points = torch.as_tensor(points, dtype=torch.float32)
    pc_range = torch.tensor(point_cloud_range).float()
    vsize = torch.tensor(voxel_size).float()
    gsize = torch.tensor(grid_size)
    voxel_x, voxel_y, voxel_z = voxel_size
    x_offset = voxel_x / 2 + point_cloud_range[0]
    y_offset = voxel_y / 2 + point_cloud_range[1]
    z_offset = voxel_z / 2 + point_cloud_range[2]
    scale_xy, scale_y = grid_size[0] * grid_size[1], grid_size[1]
    points_coords = torch.floor((points[:, [1, 2]] - pc_range[[0, 1]]) / vsize[[0, 1]]).int()
    mask = ((points_coords >= 0) & (points_coords < gsize[[0, 1]])).all(dim=1)
    points = points[mask]
    points_coords = points_coords[mask]
    points_xyz = points[:, [1, 2, 3]].contiguous()
    merge_coords = points[:, 0].int() * scale_xy + points_coords[:, 0] * scale_y + points_coords[:, 1]
    unq_coords, unq_inv, unq_cnt = torch.unique(merge_coords, return_inverse=True, return_counts=True, dim=0)
    sums = torch.zeros((unq_coords.shape[0], 3), dtype=torch.float32).index_add_(0, unq_inv, points_xyz)
    points_mean = sums / unq_cnt.float().unsqueeze(1)
    f_cluster = points_xyz - points_mean[unq_inv, :]
    f_center = torch.zeros_like(points_xyz)
    f_center[:, 0] = points_xyz[:, 0] - (points_coords[:, 0].to(points_xyz.dtype) * voxel_x + x_offset)
    f_center[:, 1] = points_xyz[:, 1] - (points_coords[:, 1].to(points_xyz.dtype) * voxel_y + y_offset)
    f_center[:, 2] = points_xyz[:, 2] - z_offset
    features = torch.cat([points[:, 1:], f_cluster, f_center], dim=-1)
    unq_coords = unq_coords.int()
    voxel_coords = torch.stack((unq_coords // scale_xy, (unq_coords % scale_xy) // scale_y, unq_coords % scale_y,
                                torch.zeros(unq_coords.shape[0]).int()), dim=1)
    voxel_coords = voxel_coords[:, [0, 3, 2, 1]]
    return features.numpy(), unq_inv.numpy(), voxel_coords.numpy()
