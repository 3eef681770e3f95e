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
