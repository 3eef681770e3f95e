"""Mirror of tools/utils.py: nms_cpu (:61-95) and post_processing (:166-233) on the GPU."""
from ..utils.postprocess import extract_boxes_onnx, nms_cpu, plot_boxes_cv2  # noqa: F401


def post_processing(img, conf_thresh, nms_thresh, output):
    """tools/utils.py:166-233 — same computation as extract_boxes_onnx (the reference duplicates it with timing prints)."""
    return extract_boxes_onnx(conf_thresh, nms_thresh, output)
