"""CPU oracle for the triton_client pre/post hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Everything under oracle/ is a CPU restatement of the reference's algorithm (niqbal996/triton_client,
mounted read-only at /root/reference while building), each function citing the reference file:line it
follows. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import it, and only as the checker or as the timed CPU baseline. triton_client_b200 never imports it.

Pinning status:
  * pre2d (incl. utils/preprocess.py::image_adjust), yolo_post, yolo_head(v4), box3d (ResidualCoder / anchors / CenterPoint
    decode, and the pcdet glue: generate_predicted_boxes, post_processing, class_agnostic_nms, multi_classes_nms, nms_gpu,
    nms_normal_gpu, CenterHead.generate_predicted_boxes, remove_low_score_nu), pillar_vfe, dynamic pillar VFE, overlay
    (RosInference._callback, plot_boxes_cv2), evaluation metrics: checked against the UNMODIFIED reference functions run in
    the build container; the generating script is oracle/make_golden.py and its outputs are committed under tests/golden/.
  * rotated IoU: checked against the reference's own iou3d_cpu.cpp compiled into oracle/_ref/. iou_normal (nms_normal_gpu)
    exists only as CUDA in the reference: restated with the one fma its nvcc build makes (read from the SASS).
  * jpeg (oracle/c/jpeg.c), cvdraw (OpenCV's line / polygon / text rasterisers): checked live against cv2 itself.
  * voxelize (spconv / det3d points_to_voxel), letterbox + YOLOv5 Detect decode (ultralytics), detectron2 dense-detector
    inference, sensor_msgs.point_cloud2.read_points, pyquaternion: the arithmetic lives in un-vendored, un-pinned third-party
    packages and the reference has no tests -> PARITY UNPINNED; restated from the published algorithms.
"""
