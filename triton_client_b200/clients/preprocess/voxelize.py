"""Drop-in for clients/preprocess/voxelize.py (det3DPreprocess, the CenterPoint / nuScenes path): zero feature
column appended, det3d VoxelGenerator semantics ('break' at max_voxels), voxels float32, batch column prepended."""
import numpy as np

from .base_preprocess import Preprocess
from .preprocess_3d import PointpillarPreprocess  # noqa: F401  (same engine plumbing)
from .. import wire

# data/nusc_centerpoint_pp_02voxel_two_pfn_10sweep.py:138-143 of the reference
_NUSC_DEFAULT = dict(range=[-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], voxel_size=[0.2, 0.2, 8], max_points_in_voxel=20, max_voxel_num=[20000, 60000])


def _voxel_generator_of(cfg_path):
    """The `voxel_generator = dict(...)` assignment of a det3d config file (plain Python; the reference reads it with
    det3d's Config.fromfile, voxelize.py:14-18). Only literals are evaluated — the file is not executed."""
    import ast
    with open(cfg_path) as f:
        tree = ast.parse(f.read(), cfg_path)
    for node in tree.body:
        if isinstance(node, ast.Assign) and any(isinstance(t, ast.Name) and t.id == "voxel_generator" for t in node.targets):
            v = node.value
            if isinstance(v, ast.Call) and getattr(v.func, "id", None) == "dict" and not v.args:
                return {k.arg: ast.literal_eval(k.value) for k in v.keywords}
            return dict(ast.literal_eval(v))
    raise ValueError("%s has no voxel_generator = dict(...) assignment" % cfg_path)


class det3DPreprocess(Preprocess):
    def __init__(self, cfg=None, engine=None, f64_math=True, **overrides):
        c = dict(_NUSC_DEFAULT)
        if cfg is not None:            # a config the caller names must be honoured or refused, never silently replaced
            got = _voxel_generator_of(cfg)
            missing = [k for k in _NUSC_DEFAULT if k not in got]
            if missing:
                raise ValueError("%s: voxel_generator lacks %s" % (cfg, missing))
            c.update({k: got[k] for k in _NUSC_DEFAULT})
        c.update(overrides)
        self._engine = engine
        self._stager = None
        self.range = c["range"]
        self.voxel_size = c["voxel_size"]
        self.max_points_in_voxel = c["max_points_in_voxel"]
        self.max_voxel_num = c["max_voxel_num"]
        # the reference hands det3d a float64 array (np.concatenate with a float64 zero column, voxelize.py:38-39), so its
        # cell indices are computed in double; f64_math=False gives the float32 arithmetic of a .bin / Triton FP32 pipeline
        self.f64_math = f64_math

    @property
    def engine(self):
        if self._engine is None:
            self._engine = wire.default_engine()
        return self._engine

    def preprocess(self):
        pass

    def filter_pc_device(self, pointcloud_array):
        import torch
        src = np.asarray(pointcloud_array)
        pts = np.ascontiguousarray(src, dtype=np.float32)
        if src.dtype == np.float64 and not np.array_equal(pts.astype(np.float64), src, equal_nan=True):
            # the reference keeps float64 points in float64 for the cell arithmetic; the kernel reads float32 rows, which is
            # the same thing only while the values are float32 numbers held in a float64 array (what read_points yields)
            raise TypeError("det3DPreprocess.filter_pc: float64 points that are not exactly float32 values are not supported")
        pts = np.concatenate((pts, np.zeros((pts.shape[0], 1), np.float32)), axis=1)      # voxelize.py:38-39
        dev = torch.from_numpy(pts).to(self.engine.device)
        return self.engine.voxelize(dev, self.range, self.voxel_size, self.max_points_in_voxel, self.max_voxel_num[0],   # [0] as the reference (:23)
                                    overflow="break", f64_math=self.f64_math)

    def filter_pc(self, pointcloud_array):
        voxels, coords, num, m = self.filter_pc_device(pointcloud_array)
        m = int(m.item())
        self.points = [voxels[:m].cpu().numpy().astype(np.float32), coords[:m].cpu().numpy(), num[:m].cpu().numpy()]
        return self.points
