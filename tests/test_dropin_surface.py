"""The drop-in package mirrors the reference's client surface (no GPU needed): names, module paths, signatures."""
import inspect

import pytest

from oracle import ref_harness


def test_names_and_module_paths():
    import triton_client_b200.clients as c
    for name in ["Yolov5client", "Client", "FCOS_client", "Pointpillars_client", "Yolov5preprocess", "FCOSpreprocess", "det3DPreprocess",
                 "PointpillarPreprocess", "Postprocess", "Yolov5postprocess", "FCOSpostprocess", "PointPillarPostprocess",
                 "YOLOv5Client", "DetectronClient", "Detector3DClient"]:
        assert hasattr(c, name), name
    import triton_client_b200.clients.preprocess.yolov5_preprocess  # noqa: F401
    import triton_client_b200.clients.postprocess.yolov5_postprocess  # noqa: F401
    import triton_client_b200.clients.postprocess.detector_3d_postprocess  # noqa: F401
    import triton_client_b200.utils.postprocess as up
    import triton_client_b200.tools.yolo_layer as yl
    for fn in ["deserialize_bytes_float", "deserialize_bytes_int", "xywh2xyxy", "box_iou", "nms_cpu", "extract_boxes_yolov5",
               "extract_boxes_onnx", "extract_boxes_triton", "extract_boxes_detectron"]:
        assert callable(getattr(up, fn))
    assert callable(yl.yolo_forward_dynamic) and hasattr(yl, "YoloLayer")


def test_wire_helpers_match_reference_values():
    import numpy as np
    from triton_client_b200.clients.postprocess.base_postprocess import Postprocess
    from oracle import yolo_post

    class P(Postprocess):
        def postprocess(self):
            pass

    buf = np.array([1.5, -2.25, 3e-8], dtype="<f4").tobytes()
    a, b = P().deserialize_bytes_float(buf), yolo_post.deserialize_bytes_float(buf)
    assert a.dtype == b.dtype == np.object_ and list(a) == list(b) and type(a[0]) is float
    ib = np.array([3, -7], dtype="<i8").tobytes()
    assert list(P().deserialize_bytes_int(ib)) == list(yolo_post.deserialize_bytes_int(ib))


@pytest.mark.skipif(not ref_harness.available(), reason="reference not mounted (GPU box)")
def test_signatures_match_live_reference():
    import triton_client_b200.clients as mine
    ref = ref_harness.clients()
    pairs = [(mine.Yolov5postprocess.extract_boxes, ref.Yolov5postprocess.extract_boxes),
             (mine.Yolov5preprocess.image_adjust, ref.Yolov5preprocess.image_adjust),
             (mine.FCOSpreprocess.image_adjust, ref.FCOSpreprocess.image_adjust),
             (mine.FCOSpostprocess.extract_boxes, ref.FCOSpostprocess.extract_boxes),
             (mine.Yolov5postprocess.load_class_names, ref.Yolov5postprocess.load_class_names),
             (mine.Postprocess.nms_cpu, ref.Postprocess.nms_cpu), (mine.Postprocess.box_iou, ref.Postprocess.box_iou),
             (mine.Postprocess.xywh2xyxy, ref.Postprocess.xywh2xyxy),
             (mine.Yolov5postprocess.plot_boxes_cv2, ref.Yolov5postprocess.plot_boxes_cv2)]
    for a, b in pairs:
        pa, pb = inspect.signature(a).parameters, inspect.signature(b).parameters
        assert list(pa) == list(pb), (a, list(pa), list(pb))
        for k in pb:
            assert pa[k].default == pb[k].default, (a, k)
    for cls in ("Yolov5client", "FCOS_client"):
        for m in ("register_client", "get_preprocess", "get_postprocess", "parse_model"):
            assert hasattr(getattr(mine, cls), m)


def test_det3d_preprocess_reads_the_config_it_is_given(tmp_path):
    """The reference builds its voxel generator from Config.fromfile(cfg) (clients/preprocess/voxelize.py:13-24): a cfg
    path must change the voxelisation parameters, and one that cannot be honoured must be refused."""
    from triton_client_b200.clients.preprocess.voxelize import det3DPreprocess
    good = tmp_path / "cfg.py"
    good.write_text("import itertools\nvoxel_generator = dict(\n    range=[-10.0, -10.0, -2.0, 10.0, 10.0, 2.0],\n    voxel_size=[0.1, 0.1, 4],\n"
                    "    max_points_in_voxel=7,\n    max_voxel_num=[111, 222],\n)\n")
    p = det3DPreprocess(str(good))
    assert p.range == [-10.0, -10.0, -2.0, 10.0, 10.0, 2.0] and p.voxel_size == [0.1, 0.1, 4]
    assert p.max_points_in_voxel == 7 and p.max_voxel_num == [111, 222]
    assert det3DPreprocess().max_points_in_voxel == 20
    bad = tmp_path / "bad.py"
    bad.write_text("voxel_generator = dict(range=[0, 0, 0, 1, 1, 1])\n")
    with pytest.raises(ValueError):
        det3DPreprocess(str(bad))
    with pytest.raises(ValueError):
        (tmp_path / "none.py").write_text("x = 1\n")
        det3DPreprocess(str(tmp_path / "none.py"))
