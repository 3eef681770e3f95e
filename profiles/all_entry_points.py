#!/usr/bin/env python
"""Small end-to-end exercise of every C-ABI entry point, checked against the oracle.

Written as the workload for `compute-sanitizer --tool memcheck|racecheck`; on this pool
compute-sanitizer is closed ("runs under it have left GPUs needing a reset"), so it is run plain as a
fast all-entry-point check:  python profiles/all_entry_points.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import box3d, detectron_post, pre2d, yolo_post  # noqa: E402
from oracle import voxelize as ov  # noqa: E402
from triton_client_b200 import gpu, synth  # noqa: E402

eng = gpu.Engine(0, workspace_bytes=256 << 20)
c = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731

for (h, w, oh, ow, mode) in [(90, 160, 64, 64, "stretch"), (180, 320, 96, 160, "letterbox"), (33, 47, 50, 61, "stretch"), (40, 40, 48, 56, "pad")]:
    img = synth.image_noise(h, w, 1)
    pv = 0 if mode == "pad" else 114
    out, _ = eng.preprocess_images(c(img[None]), oh, ow, mode=mode, pad_value=pv)
    assert np.array_equal(out.cpu().numpy(), pre2d.preprocess_frame(img, oh, ow, mode=mode, pad_value=pv)), mode
big = synth.image_noise(200, 300, 2)
view = c(big)[7:150, 5:222]   # unaligned strided view: exercises the buffer-edge path
out, _ = eng.preprocess_images(view, 80, 96)
assert np.array_equal(out.cpu().numpy(), pre2d.preprocess_frame(view.cpu().numpy(), 80, 96))

for (b, n, nc) in [(2, 3000, 2), (1, 2000, 80), (1, 9000, 3)]:
    pred = synth.yolo_output(b, n, nc, 5, objects=20, rows_per_object=20)
    if n == 9000:   # > 4096 candidates: ordering in global scratch, streaming NMS rounds
        pred[..., 4] = np.maximum(pred[..., 4], 0.5)
        pred[..., 5:] = np.maximum(pred[..., 5:], 0.7)
    o, cnt, _ = eng.yolo_nms(c(pred), conf_thres=0.3)
    ref = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False)
    for i in range(b):
        assert torch.equal(o[i, :cnt[i]].cpu(), ref[i].double().reshape(-1, 6))
    gpu.yolo_nms_static(c(pred), conf_thres=0.3)

pts = synth.lidar_sweep(9, beams=16, azimuths=400)
K = ([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4])
for ovf, mv in (("continue", 40000), ("break", 300)):
    v, co, nn, m = eng.voxelize(c(pts), K[0], K[1], 32, mv, overflow=ovf)
    mm = int(m.item())
    want = ov.points_to_voxel(pts, K[0], K[1], 32, mv, ovf)
    assert np.array_equal(v[:mm].cpu().numpy(), want[0]) and np.array_equal(nn[:mm].cpu().numpy(), want[2])
eng.pillar_decorate(v, co, nn, K[1], K[0], m=m)
eng.dynamic_pillar(c(np.concatenate([np.zeros((len(pts), 1), np.float32), pts], 1)), [-51.2, -51.2, -5, 51.2, 51.2, 3], [0.2, 0.2, 8])

PP = dict(anchor_range=[0, -39.68, -3, 69.12, 39.68, 1], anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]],
          anchor_bottoms=[-1.78, -0.6, -0.6], anchor_rotations=[[0, 1.57]] * 3)
cls, box, dirp = synth.pp_head_outputs(3, h=40, w=36, planted=10)
ob, osc, ol, cnt = eng.pp_anchor_post(c(cls), c(box), c(dirp), **PP)
ref = box3d.pp_anchor_post(cls, box, dirp, **PP)
assert np.array_equal(ol[:int(cnt)].cpu().numpy(), ref["pred_labels"].numpy())
heads = synth.centerpoint_head_outputs(4, classes_per_head=(1, 2), h=32, w=32, peaks=10)
eng.centerpoint_post([{k: c(t) for k, t in hd.items()} for hd in heads], [[0], [1, 2]], 4, [0.2, 0.2, 8], [-51.2, -51.2],
                     [-61.2, -61.2, -10, 61.2, 61.2, 10], k=100)
bx = synth.random_boxes7(600, 3, extent=30)
sc = np.random.default_rng(1).uniform(0, 1, 600).astype(np.float32)
keep, cnt = eng.rotated_nms(c(bx), c(sc), 0.1)
assert keep[:int(cnt)].cpu().tolist() == box3d.nms_gpu(bx, sc, 0.1)[:500].tolist()
eng.rotated_iou(c(bx[:50]), c(bx[50:120]))
lab = np.random.default_rng(2).integers(0, 10, 100)
eng.class_score_filter(c(lab.astype(np.int64)), c(sc[:100]), box3d.NUSC_GROUPS)

anchors = [12, 16, 19, 36, 40, 28]
head = np.random.default_rng(3).normal(0, 1.5, (1, 3 * 9, 8, 8)).astype(np.float32)
b4, c4 = eng.yolo_head_decode_v4([c(head)], [[a / 8 for a in anchors]], 4)
eng.yolo_head_decode_v5([c(head)], [anchors], [8], 4)
eng.yolo4_post(b4, c4, 0.1, 0.5)
eng.nms_axis_aligned(c(np.abs(bx[:, :4])), c(sc), 0.5)
eng.box_iou(c(np.abs(bx[:64, :4])), c(np.abs(bx[64:128, :4])))
lg, dl = synth.detectron_head_outputs(6, num_classes=4, image_hw=(128, 160), planted=40)
eng.detectron_post([c(t) for t in lg], [c(t) for t in dl], (8, 16, 32, 64, 128), detectron_post.retinanet_cell_anchors(), (128, 160))
from oracle import overlay  # noqa: E402
fr = synth.image_noise(120, 200, 6)
dd = np.array([[10, 40, 60, 80, 0.875, 1], [-5, 20, 30, 50, 0.5, 0]], np.float64)
got = eng.draw_detections(c(fr), c(dd[None]), None, ["crop", "weed"], (0, 255, 0), [100, 100], math_mode=1)
assert np.array_equal(got.cpu().numpy(), overlay.draw_detections(fr.copy(), dd, ["crop", "weed"], (0, 255, 0), [100, 100]))
assert np.array_equal(eng.scale_boxes(c(dd), (120, 200), (100, 100)).cpu().numpy(), overlay.scale_box_array(dd, (120, 200), (100, 100)))
from oracle import pointcloud2 as opc2  # noqa: E402
sw = synth.lidar_sweep(3, beams=8, azimuths=100)[:640]
sw[::11, 1] = np.nan
flds = [opc2.make_field(nm, 4 * k, 7) for k, nm in enumerate(("x", "y", "z", "intensity"))]
cloud = opc2.make_cloud(dict(x=sw[:, 0], y=sw[:, 1], z=sw[:, 2], intensity=sw[:, 3]), flds, 20, height=4, row_pad=5)
rows_, cnt_ = eng.pointcloud2_unpack(c(np.frombuffer(cloud.data, np.uint8)), cloud.width, 4, 20, cloud.row_step, [0, 4, 8, 12], [7] * 4)
want_ = np.array(list(opc2.read_points(cloud, skip_nans=True)))
assert int(cnt_) == len(want_) and np.array_equal(rows_[:int(cnt_)].cpu().numpy().astype(np.float64), want_)
jr, ji, jc = eng.boxes_to_jsk(c(bx[:100]), c(sc[:100]), c(lab % 3 + 1))
assert ji[:int(jc)].cpu().tolist() == opc2.jsk_rows(bx[:100], sc[:100], lab % 3 + 1)[0]
torch.cuda.synchronize()
print(f"all_entry_points: every entry point ran and matched the oracle ({eng.launches} kernel launches)")
