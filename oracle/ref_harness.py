"""Imports the UNMODIFIED reference from /root/reference (build container only — that path does not
exist on the GPU box). Used by oracle/make_golden.py to produce the fixtures under tests/golden/ and by
the `-m "not gpu"` tests that validate the oracle restatements when the reference is mounted.
Test infrastructure only; recipe from SURVEY.md Appendix B.
"""
import importlib.util
import os
import sys
import types

REF = os.environ.get("TCB_REFERENCE_ROOT", "/root/reference")
OPCD = os.path.join(REF, "utils", "export_models", "OpenPCDet")


def available():
    return os.path.isdir(os.path.join(REF, "clients"))


def _stub(name):
    if name not in sys.modules:
        m = types.ModuleType(name)
        sys.modules[name] = m
        if "." in name:
            parent, child = name.rsplit(".", 1)
            _stub(parent)
            setattr(sys.modules[parent], child, m)
    return sys.modules[name]


_clients = None


def clients():
    """The reference's `clients` package (Yolov5client, FCOS_client, pre/postprocess classes)."""
    global _clients
    if _clients is None:
        for n in ["tritonclient", "tritonclient.grpc", "tritonclient.grpc.model_config_pb2",
                  "tritonclient.grpc.service_pb2", "tritonclient.grpc.service_pb2_grpc"]:
            _stub(n)
        for n in ["jsk_recognition_msgs", "jsk_recognition_msgs.msg"]:      # ROS message classes: named, never instantiated here
            if n not in sys.modules:
                sys.modules[n] = _AnyModule(n)
        saved_path, saved_cwd = list(sys.path), os.getcwd()
        saved_mods = {k: v for k, v in sys.modules.items() if k == "clients" or k.startswith("clients.")}
        for k in saved_mods:
            del sys.modules[k]
        sys.path.insert(0, REF)
        os.chdir(REF)  # class-name files are opened by relative path (yolov5_postprocess.py:19)
        try:
            import clients as c  # noqa
            _clients = c
        finally:
            os.chdir(saved_cwd)
            sys.path[:] = saved_path
    return _clients


def load_file(modname, relpath, stubs=()):
    """Load one reference source file as a module (for pcdet pieces that cannot be imported as a package)."""
    for s in stubs:
        _stub(s)
    path = os.path.join(REF, relpath)
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def utils_postprocess():
    return load_file("ref_utils_postprocess", "utils/postprocess.py")


def yolo_layer():
    saved = list(sys.path)
    sys.path.insert(0, REF)
    try:
        return load_file("ref_yolo_layer", "tools/yolo_layer.py")
    finally:
        sys.path[:] = saved


def tools_utils():
    return load_file("ref_tools_utils", "tools/utils.py")


def box_coder_utils():
    return load_file("ref_box_coder_utils", "utils/export_models/OpenPCDet/pcdet/utils/box_coder_utils.py")


def centernet_utils():
    return load_file("ref_centernet_utils", "utils/export_models/OpenPCDet/pcdet/models/model_utils/centernet_utils.py")


def common_utils():
    return load_file("ref_common_utils", "utils/export_models/OpenPCDet/pcdet/utils/common_utils.py", stubs=("SharedArray",))


def load_package(pkgname, reldir, files):
    """Load a few sibling files of the reference as a synthetic package so their relative imports work."""
    pkg = types.ModuleType(pkgname)
    pkg.__path__ = [os.path.join(REF, reldir)]
    sys.modules[pkgname] = pkg
    mods = {}
    for f in files:
        name = f"{pkgname}.{f[:-3]}"
        spec = importlib.util.spec_from_file_location(name, os.path.join(REF, reldir, f))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        setattr(pkg, f[:-3], mod)
        mods[f[:-3]] = mod
    return mods


class _Cfg(dict):
    """EasyDict stand-in (attribute access + .get)."""
    __getattr__ = dict.__getitem__


def pillar_vfe(voxel_size, point_cloud_range, num_point_features=4):
    """The reference's PillarVFE (OPCD/pcdet/models/backbones_3d/vfe/pillar_vfe.py) with the PFN layers
    removed, so forward() returns the decorated (M,P,F+6) features of lines :94-118."""
    import torch
    mods = load_package("ref_vfe", "utils/export_models/OpenPCDet/pcdet/models/backbones_3d/vfe", ["vfe_template.py", "pillar_vfe.py"])
    cfg = _Cfg(USE_NORM=True, WITH_DISTANCE=False, USE_ABSLOTE_XYZ=True, NUM_FILTERS=[64])
    vfe = mods["pillar_vfe"].PillarVFE(cfg, num_point_features, voxel_size, point_cloud_range)
    vfe.pfn_layers = torch.nn.ModuleList([])
    return vfe


class _AnyModule(types.ModuleType):
    """Stand-in for a ROS package: any attribute is a placeholder class (so `from x.msg import Image` works)."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        cls = type(name, (), {})
        setattr(self, name, cls)
        return cls


_communicator = None


def communicator():
    """The reference's `communicator` package (RosInference._callback / _scale_boxes, EvaluateInference) with the
    ROS middleware stubbed out: rospy, sensor_msgs, vision_msgs, cv_bridge are placeholders, never called by the
    pieces the golden scripts run (those get fake bridge / channel / publisher objects)."""
    global _communicator
    if _communicator is None:
        clients()
        for n in ["rospy", "rosbag", "sensor_msgs", "sensor_msgs.msg", "vision_msgs", "vision_msgs.msg", "cv_bridge",
                  "geometry_msgs", "geometry_msgs.msg", "ros_numpy"]:
            if n not in sys.modules:
                m = _AnyModule(n)
                sys.modules[n] = m
                if "." in n:
                    setattr(sys.modules[n.rsplit(".", 1)[0]], n.rsplit(".", 1)[1], m)
        for n in ["service_pb2", "service_pb2_grpc", "model_config_pb2"]:
            setattr(sys.modules["tritonclient.grpc"], n, sys.modules["tritonclient.grpc." + n])
        saved_path, saved_cwd = list(sys.path), os.getcwd()
        sys.path.insert(0, REF)
        os.chdir(REF)
        try:
            import communicator as c  # noqa
            _communicator = c
        finally:
            os.chdir(saved_cwd)
            sys.path[:] = saved_path
    return _communicator


# ---- the vendored pcdet tree as an importable package --------------------------------------------------------------
# `import pcdet` fails here (its __init__ wants a generated version.py and every sub-package __init__ imports compiled
# ops), so the tree is mounted under the name `ref_pcdet` by a finder with three rules: a name that is a .py file of
# the tree is executed UNMODIFIED; a name that is a directory becomes an empty package (its __init__.py is not run);
# any other name below ref_pcdet (compiled extensions such as iou3d_nms_cuda, generated files) is a placeholder module.
# Third-party packages the tree imports and this image lacks (spconv, torch_scatter, SharedArray, ...) are placeholders
# too, except where a golden script installs a working stand-in first (torch_scatter, iou3d_nms_cuda below).
import importlib.abc
import importlib.machinery

PCDET = "ref_pcdet"


class _PcdetFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname != PCDET and not fullname.startswith(PCDET + "."):
            return None
        rel = fullname.split(".")[1:]
        base = os.path.join(OPCD, "pcdet", *rel)
        if os.path.isfile(base + ".py"):
            return importlib.util.spec_from_file_location(fullname, base + ".py")
        if os.path.isdir(base):
            spec = importlib.machinery.ModuleSpec(fullname, self, is_package=True)
            spec.submodule_search_locations = [base]
            return spec
        return importlib.machinery.ModuleSpec(fullname, self)

    def create_module(self, spec):
        if spec.submodule_search_locations is None:
            return _AnyModule(spec.name)
        return None

    def exec_module(self, module):
        pass


def pcdet(module):
    """Import `pcdet.<module>` of the vendored OpenPCDet tree, e.g. pcdet("models.model_utils.model_nms_utils")."""
    if not any(isinstance(f, _PcdetFinder) for f in sys.meta_path):
        sys.meta_path.insert(0, _PcdetFinder())
        for n in ["SharedArray", "spconv", "spconv.pytorch", "cumm", "easydict", "tensorboardX", "numba", "kornia"]:
            try:
                importlib.import_module(n)
            except Exception:
                if n not in sys.modules:
                    sys.modules[n] = _AnyModule(n)
    return importlib.import_module(PCDET + "." + module)


class no_cuda:
    """Tensor.cuda() / Module.cuda() as no-ops while reference code that hard-codes `.cuda()` runs (no GPU here)."""

    def __enter__(self):
        import torch
        self._t, self._m = torch.Tensor.cuda, torch.nn.Module.cuda
        torch.Tensor.cuda = lambda s, *a, **k: s
        torch.nn.Module.cuda = lambda s, *a, **k: s

    def __exit__(self, *exc):
        import torch
        torch.Tensor.cuda, torch.nn.Module.cuda = self._t, self._m


def install_torch_scatter():
    """torch_scatter is not in this image. Stand-in with its published semantics for the two calls the tree makes
    (dynamic_pillar_vfe.py:40,105): scatter_mean = scatter_add in input order / count (count clamped to 1),
    scatter_max = per-group maximum and arg-max."""
    import torch
    if "torch_scatter" in sys.modules and hasattr(sys.modules["torch_scatter"], "_tcb_standin"):
        return
    m = types.ModuleType("torch_scatter")
    m._tcb_standin = True

    def scatter_mean(src, index, dim=0):
        assert dim == 0
        n = int(index.max()) + 1 if index.numel() else 0
        out = torch.zeros((n,) + tuple(src.shape[1:]), dtype=src.dtype).index_add_(0, index, src)
        cnt = torch.zeros((n,), dtype=src.dtype).index_add_(0, index, torch.ones_like(index, dtype=src.dtype)).clamp_(min=1)
        return out / cnt.view((n,) + (1,) * (src.dim() - 1))

    def scatter_max(src, index, dim=0):
        assert dim == 0
        n = int(index.max()) + 1 if index.numel() else 0
        idx = index.view((-1,) + (1,) * (src.dim() - 1)).expand_as(src)
        out = torch.full((n,) + tuple(src.shape[1:]), float("-inf"), dtype=src.dtype).scatter_reduce_(0, idx, src, "amax", include_self=True)
        return out, None

    m.scatter_mean, m.scatter_max = scatter_mean, scatter_max
    sys.modules["torch_scatter"] = m


def install_iou3d_nms_cuda():
    """The compiled extension behind pcdet.ops.iou3d_nms.iou3d_nms_utils, rebuilt for the CPU from the reference's own
    parts: the pairwise test is the reference's iou_bev as compiled from its iou3d_cpu.cpp (oracle/_ref/libiou3d_ref.so;
    same arithmetic as the device function, iou3d_nms_kernel.cu:227-234), the mask layout and the host sweep are
    iou3d_nms.cpp:90-136 (box i suppresses a later box j when iou_bev(i, j) > thresh). nms_normal_gpu uses the
    axis-aligned overlap of iou3d_nms_kernel.cu:314-325, which has no CPU twin in the reference: box3d.iou_normal restates
    it with the one contraction the reference's own nvcc build makes."""
    import ctypes
    import numpy as np
    import torch
    name = PCDET + ".ops.iou3d_nms.iou3d_nms_cuda"
    if name in sys.modules and hasattr(sys.modules[name], "_tcb_standin"):
        return
    pcdet("ops.iou3d_nms")    # the package first
    so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libiou3d_ref.so")
    ref = ctypes.CDLL(so)
    m = types.ModuleType(name)
    m._tcb_standin = True

    def sweep(over, keep):
        n = over.shape[0]
        removed = np.zeros(n, bool)
        k = 0
        for i in range(n):
            if not removed[i]:
                keep[k] = i
                k += 1
                removed[i + 1:] |= over[i, i + 1:]
        return k

    def nms_gpu(boxes, keep, thresh):
        b = np.ascontiguousarray(boxes.detach().numpy(), np.float32)
        n = b.shape[0]
        iou = np.zeros((n, n), np.float32)
        P = ctypes.c_void_p
        if n:
            ref.ref_boxes_iou_bev_cpu(P(b.ctypes.data), n, P(b.ctypes.data), n, P(iou.ctypes.data))
        return sweep(iou > np.float32(thresh), keep)

    def nms_normal_gpu(boxes, keep, thresh):
        from oracle import box3d
        b = np.ascontiguousarray(boxes.detach().numpy(), np.float32)
        return sweep(box3d.iou_normal(b) > np.float32(thresh), keep)

    m.nms_gpu, m.nms_normal_gpu = nms_gpu, nms_normal_gpu
    sys.modules[name] = m
    setattr(sys.modules[PCDET + ".ops.iou3d_nms"], "iou3d_nms_cuda", m)
