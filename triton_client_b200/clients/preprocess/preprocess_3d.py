"""Drop-in for clients/preprocess/preprocess_3d.py (PointpillarPreprocess): same constructor behaviour (reads
./data/kitti_dataset.yaml when present), same filter_pc(points) -> dict contract, voxelisation on the GPU.

The reference builds pcdet's PointFeatureEncoder + DataProcessor (preprocess_3d.py:12-27) whose test-time chain is:
absolute_coordinates_encoding (x,y,z,intensity kept) -> mask_points_by_range (x/y inclusive) -> no shuffle ->
spconv VoxelGenerator ('continue' at max_voxels) -> batch column prepended (:50)."""
import os

import numpy as np
import torch

from .base_preprocess import Preprocess
from .. import wire

# data/kitti_dataset.yaml:1,64-70 of the reference (the config the shipped client loads)
_KITTI_DEFAULT = dict(POINT_CLOUD_RANGE=[0, -40, -3, 70.4, 40, 1], VOXEL_SIZE=[0.05, 0.05, 0.1], MAX_POINTS_PER_VOXEL=5,
                      MAX_NUMBER_OF_VOXELS={"train": 16000, "test": 40000}, NUM_POINT_FEATURES=4)


def _load_dataset_cfg(path):
    cfg = dict(_KITTI_DEFAULT)
    if path and os.path.exists(path):
        import yaml
        with open(path, "r") as f:
            y = yaml.safe_load(f)
        cfg["POINT_CLOUD_RANGE"] = y.get("POINT_CLOUD_RANGE", cfg["POINT_CLOUD_RANGE"])
        for proc in y.get("DATA_PROCESSOR", []):
            if proc.get("NAME") == "transform_points_to_voxels":
                cfg["VOXEL_SIZE"] = proc["VOXEL_SIZE"]
                cfg["MAX_POINTS_PER_VOXEL"] = proc["MAX_POINTS_PER_VOXEL"]
                cfg["MAX_NUMBER_OF_VOXELS"] = proc["MAX_NUMBER_OF_VOXELS"]
        used = y.get("POINT_FEATURE_ENCODING", {}).get("used_feature_list")
        if used:
            cfg["NUM_POINT_FEATURES"] = len(used)
    return cfg


class PointpillarPreprocess(Preprocess):
    def __init__(self, engine=None, cfg_path='./data/kitti_dataset.yaml', **overrides):
        self._engine = engine
        self._stager = None
        self.dataset_cfg = _load_dataset_cfg(cfg_path)
        self.dataset_cfg.update(overrides)
        self.point_cloud_range = np.array(self.dataset_cfg["POINT_CLOUD_RANGE"], dtype=np.float32)
        self.voxel_size = self.dataset_cfg["VOXEL_SIZE"]
        self.max_points = int(self.dataset_cfg["MAX_POINTS_PER_VOXEL"])
        self.max_voxels = int(self.dataset_cfg["MAX_NUMBER_OF_VOXELS"]["test"])   # training=False (preprocess_3d.py:24)
        self.num_point_features = int(self.dataset_cfg["NUM_POINT_FEATURES"])

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

    def filter_pc_device(self, pointcloud_array, z_offset=0.0, intensity_col=-1):
        """Voxelise on the device; returns CUDA tensors (voxels, voxel_coords [b,z,y,x], voxel_num_points, m).
        z_offset / intensity_col fuse the caller's conditioning (communicator/ros_inference3d.py:127-128)."""
        pts = np.ascontiguousarray(pointcloud_array, dtype=np.float32) if isinstance(pointcloud_array, np.ndarray) else pointcloud_array
        dev = self.stager.to_device(pts, np.float32, tuple(pts.shape), "pc_in")
        return self.engine.voxelize(dev, [float(v) for v in self.point_cloud_range], self.voxel_size, self.max_points, self.max_voxels,
                                    overflow="continue", range_mask=True, z_offset=z_offset, intensity_col=intensity_col,
                                    nfeat=self.num_point_features)

    def filter_pc(self, pointcloud_array):
        '''
        pointcloud_array: input numpy array of shape [N,3] or [N,4] with/without reflectivity
        return: Dictionary with following keys:
        data_dict['points'] = (N', 4) points inside the x/y range
        data_dict['voxels'] = (M, MAX_POINTS_PER_VOXEL, 4) float32
        data_dict['voxel_coords'] = (M, 4) int32 [batch, z, y, x]
        data_dict['voxel_num_points'] = (M,) int32
        '''
        pts = np.asarray(pointcloud_array)
        voxels, coords, num, m = self.filter_pc_device(pts)
        m = int(m.item())
        r = self.point_cloud_range
        mask = (pts[:, 0] >= r[0]) & (pts[:, 0] <= r[3]) & (pts[:, 1] >= r[1]) & (pts[:, 1] <= r[4])   # bookkeeping copy of the kept points
        return {"points": pts[mask], "frame_id": 1, "use_lead_xyz": True,
                "voxels": voxels[:m].cpu().numpy(), "voxel_coords": coords[:m].cpu().numpy(), "voxel_num_points": num[:m].cpu().numpy()}
