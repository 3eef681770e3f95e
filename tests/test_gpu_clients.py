"""The drop-in client classes on the GPU: same results and conventions as the reference's classes (golden vectors,
oracle), including the return-the-exception convention and object/float64 output contracts."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import box3d, pre2d, yolo_head, yolo_post
from oracle import voxelize as ov
from triton_client_b200 import synth
from triton_client_b200.clients import (FCOS_client, HostResponse, PointPillarPostprocess, PointpillarPreprocess, Pointpillars_client,
                                        Yolov5client, det3DPreprocess)

pytestmark = pytest.mark.gpu


def test_yolov5_client_golden(engine):
    client = Yolov5client(engine=engine)
    pre, post = client.get_preprocess(), client.get_postprocess()
    g = golden("pre2d")
    import cv2
    for k in range(5):
        h, w, oh, ow = (int(v) for v in g[f"shape{k}"])
        img = g[f"img{k}"]
        resized = cv2.resize(img, (ow, oh))
        before = resized.copy()
        out = pre.image_adjust(resized)                      # the reference call: already-resized RGB frame
        assert out.dtype == np.float32 and out.shape == (1, 3, oh, ow) and np.array_equal(out, g[f"yolo{k}"])
        assert np.array_equal(resized, before)               # input not mutated
        fused, aff = pre.resize_adjust(img, (oh, ow))        # added: resize fused in
        assert np.array_equal(fused, g[f"yolo{k}"])
    gy = golden("yolo_post")
    for k in range(4):
        conf, iou, agn, max_det = gy[f"cfg{k}"]
        pred = gy[f"pred{k}"]
        res = post.extract_boxes(HostResponse([pred.tobytes()], [pred.shape]), conf_thres=float(conf), iou_thres=float(iou),
                                 classes=gy[f"classes{k}"].tolist() or None, agnostic=bool(agn), max_det=int(max_det))
        assert isinstance(res, list) and len(res) == pred.shape[0]
        for b, t in enumerate(res):
            want = gy[f"out{k}_{b}"].reshape(-1, 6)
            assert t.shape == want.shape and (t.shape[0] == 0 or t.dtype == torch.float64)
            assert np.array_equal(t.numpy().reshape(-1, 6), want)


def test_yolov5_error_convention(engine):
    post = Yolov5client(engine=engine).get_postprocess()
    bg = np.zeros((1, 64, 7), np.float32)
    r = post.extract_boxes(HostResponse([bg]), conf_thres=0.3)
    assert isinstance(r, AssertionError) and str(r) == str(golden("yolo_post")["empty_msg"])
    pred = synth.yolo_output(1, 500, 2, 1)
    r = post.extract_boxes(HostResponse([pred]), conf_thres=1.5)
    assert isinstance(r, AssertionError) and "Invalid Confidence threshold" in str(r) or "No valid detections" in str(r)
    r = post.extract_boxes(HostResponse([pred]), conf_thres=0.3, iou_thres=2.0)
    assert isinstance(r, AssertionError) and "Invalid IoU" in str(r)
    assert isinstance(post.extract_boxes(HostResponse([pred]), conf_thres=0.3, multi_label=True), TypeError)
    assert isinstance(post.extract_boxes(object()), Exception)          # malformed response: returned, not raised
    ok = post.extract_boxes(HostResponse([pred]), conf_thres=0.3)
    assert isinstance(ok, list)


def test_fcos_client(engine):
    c = FCOS_client(engine=engine)
    img = synth.image_noise(50, 70, 3)
    assert np.array_equal(c.get_preprocess().image_adjust(img), pre2d.image_adjust_fcos(img))
    boxes = np.random.default_rng(0).uniform(0, 100, (7, 4)).astype(np.float32)
    classes = np.arange(7, dtype=np.int64)
    scores = np.linspace(0.1, 0.9, 7).astype(np.float32)
    b, ci, s = c.get_postprocess().extract_boxes(HostResponse([boxes.tobytes(), classes.tobytes(), scores.tobytes()], [(7, 4), (7,), (7,)]))
    assert b.shape == (7, 4) and b.dtype == np.object_ and list(ci) == list(range(7)) and float(s[3]) == float(scores[3])


def test_pointpillars_client_preprocess(engine):
    client = Pointpillars_client(engine=engine)
    pre = client.get_preprocess()
    assert isinstance(pre, PointpillarPreprocess)
    pts = synth.lidar_sweep(synth.seed_for(4, 7))
    d = pre.filter_pc(pts)
    want = ov.filter_pc_pointpillar(pts, pre.point_cloud_range, pre.voxel_size, pre.max_points, pre.max_voxels)
    assert d["voxels"].dtype == np.float32 and d["voxel_coords"].dtype == np.int32 and d["voxel_num_points"].dtype == np.int32
    for k in ("voxels", "voxel_coords", "voxel_num_points", "points"):
        assert np.array_equal(d[k], want[k]), k
    pillar = PointpillarPreprocess(engine, cfg_path=None, POINT_CLOUD_RANGE=[0, -39.68, -3, 69.12, 39.68, 1], VOXEL_SIZE=[0.16, 0.16, 4],
                                   MAX_POINTS_PER_VOXEL=32, MAX_NUMBER_OF_VOXELS={"test": 10000})
    d = pillar.filter_pc(pts)
    want = ov.filter_pc_pointpillar(pts, [0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4], 32, 10000)
    assert d["voxels"].shape[0] == 10000 and all(np.array_equal(d[k], want[k]) for k in ("voxels", "voxel_coords", "voxel_num_points"))


def test_det3d_preprocess(engine):
    pts = synth.lidar_multisweep(synth.seed_for(5, 3))[:, :4]
    for f64 in (True, False):
        got = det3DPreprocess(engine=engine, f64_math=f64).filter_pc(pts)
        want = ov.filter_pc_det3d(pts, [-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8], 20, 20000, f64_math=f64)
        assert got[0].dtype == np.float32 and got[0].shape[2] == 5 and got[1].shape[1] == 4
        assert all(np.array_equal(a, b) for a, b in zip(got, want))


def test_pointpillar_postprocess(engine):
    post = Pointpillars_client(engine=engine).get_postprocess()
    assert isinstance(post, PointPillarPostprocess)
    rng = np.random.default_rng(2)
    boxes = rng.normal(0, 10, (40, 7)).astype(np.float32)
    scores = (np.round(rng.uniform(0, 1, 40) * 20) / 20).astype(np.float32)
    labels = rng.integers(0, 10, 40).astype(np.int64)
    out = post.extract_boxes(HostResponse([boxes.tobytes(), scores.tobytes(), labels.tobytes()], [(40, 7), (40,), (40,)]))
    assert out["pred_boxes"].shape == (40, 7) and out["pred_boxes"].dtype == np.object_ and int(out["pred_labels"][5]) == int(labels[5])
    pred = {"labels": labels, "scores": scores, "box3d_lidar": boxes}
    got = post.remove_low_score_nu(pred, 0.1)
    want = box3d.remove_low_score_nu(pred)
    assert all(np.array_equal(got[k], want[k]) for k in pred)
    assert post.get_annotations_indices(8, 0.1, labels, scores) == box3d.get_annotations_indices(8, 0.1, labels, scores)
    cls, box, dirp = synth.pp_head_outputs(4010, h=60, w=56, planted=12)
    res = post.decode_pointpillars(cls, box, dirp)
    ref = box3d.pp_anchor_post(cls, box, dirp, anchor_range=[0, -39.68, -3, 69.12, 39.68, 1],
                               anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]], anchor_bottoms=[-1.78, -0.6, -0.6],
                               anchor_rotations=[[0, 1.57]] * 3)
    assert torch.equal(res["pred_labels"], ref["pred_labels"])
    np.testing.assert_allclose(res["pred_boxes"].numpy(), ref["pred_boxes"].numpy(), rtol=1e-5, atol=1e-5)
    heads = synth.centerpoint_head_outputs(5010)
    res = post.decode_centerpoint(heads)
    ref = box3d.centerhead_generate_predicted_boxes(heads, [[0], [1, 2], [3, 4], [5], [6, 7], [8, 9]], [-51.2, -51.2, -5.0, 51.2, 51.2, 3.0],
                                                    [0.2, 0.2, 8], 4, [-61.2, -61.2, -10.0, 61.2, 61.2, 10.0])
    assert torch.equal(res["pred_labels"], ref["pred_labels"])


def test_utils_and_tools_twins(engine):
    from triton_client_b200.clients import wire
    wire._engine = engine
    from triton_client_b200.tools import yolo_layer as yl
    from triton_client_b200.utils import postprocess as up
    from triton_client_b200.utils import preprocess as upre
    g = golden("yolo4")
    layer = yl.YoloLayer(anchor_mask=[0, 1, 2], num_classes=4, anchors=g["anchors"].tolist(), num_anchors=9, stride=8)
    layer.eval()
    b, c = layer(torch.from_numpy(g["head0"]))
    np.testing.assert_allclose(b.cpu().numpy(), g["boxes0"], rtol=1e-5, atol=1e-6)
    boxes = np.concatenate([g[f"boxes{i}"] for i in range(3)], 1)
    confs = np.concatenate([g[f"confs{i}"] for i in range(3)], 1)
    rows = up.extract_boxes_onnx(0.2, 0.4, [boxes, confs])
    for i, r in enumerate(rows):
        assert np.array_equal(np.array(r, np.float32).reshape(-1, 7), g[f"post{i}"])
    pred = synth.yolo_output(1, 400, 3, 9)
    want = yolo_post.extract_boxes(pred, conf_thres=0.3, as_written=False)
    got = up.extract_boxes_yolov5(HostResponse([pred]), conf_thres=0.3)
    assert torch.equal(got[0], want[0])
    with pytest.raises(AssertionError):
        up.extract_boxes_yolov5(HostResponse([np.zeros((1, 8, 7), np.float32)]), conf_thres=0.3)
    bgr = synth.image_noise(60, 80, 4)
    assert np.array_equal(upre.image_adjust(bgr, 32, 24, "COCO"), pre2d.image_adjust_utils(bgr, 32, 24, "COCO"))
    assert np.array_equal(upre.image_adjust_ros(bgr), pre2d.image_adjust_ros(bgr))
    bx = np.random.default_rng(1).uniform(0, 50, (30, 2)).astype(np.float32)
    bx = np.concatenate([bx, bx + 10], 1)
    sc = np.random.default_rng(2).permutation(30).astype(np.float32)
    assert np.array_equal(up.nms_cpu(bx, sc, 0.3), yolo_post.nms_cpu(bx, sc, 0.3))
    assert torch.equal(up.box_iou(torch.from_numpy(bx), torch.from_numpy(bx[:5])), yolo_post.box_iou(torch.from_numpy(bx), torch.from_numpy(bx[:5])))


def test_pipelined_batch_equals_single_shot(engine):
    """image_adjust_batch cuts big batches into chunks over three streams; results must not depend on that."""
    pre = Yolov5client(engine=engine).get_preprocess()
    frames = np.stack([synth.image_noise(180, 320, 900 + i) for i in range(19)])          # 19: ragged last chunk
    a = pre.image_adjust_batch(frames, (160, 160), mode="letterbox", chunks=4)
    b = pre.image_adjust_batch(torch.from_numpy(frames).pin_memory(), (160, 160), mode="letterbox", chunks=1)
    c = pre.image_adjust_batch(frames, (160, 160), mode="letterbox", chunks=4, copy=False)
    want = np.concatenate([pre2d.preprocess_frame(f, 160, 160, mode="letterbox") for f in frames])
    assert np.array_equal(a, want) and np.array_equal(b, want) and np.array_equal(c, want)
    a2 = pre.image_adjust_batch(frames[::-1].copy(), (160, 160), mode="letterbox", chunks=4)   # buffers reused across calls
    assert np.array_equal(a2, want[::-1])


@pytest.mark.parametrize("hw", [(180, 320), (320, 180), (200, 200)])     # bands above/below, bands left/right, no bands
def test_repeated_batches_download_only_the_frame_rows(engine, hw):
    """From the second batch of a geometry on, only the rectangle that carries the frame comes back over PCIe (the padding
    already sits in the pinned result buffer): results stay those of a full download, also when the geometry changes in
    between and when deferred calls alternate between the two result buffers."""
    pre = Yolov5client(engine=engine).get_preprocess()
    h, w = hw
    mk = lambda seed, hh=h, ww=w, n=12: np.stack([synth.image_noise(hh, ww, seed + i) for i in range(n)])
    want = lambda fr: np.concatenate([pre2d.preprocess_frame(f, 160, 160, mode="letterbox") for f in fr])
    sizes = []
    for k in range(4):
        fr = mk(3000 + 20 * k)
        got = pre.image_adjust_batch(fr, (160, 160), mode="letterbox", chunks=3, copy=(k % 2 == 0))
        assert np.array_equal(got, want(fr)), k
        sizes.append(pre.last_d2h_bytes)
    full = 12 * 3 * 160 * 160 * 4
    assert sizes[0] == full and (sizes[1] < full if h != w else sizes[1] == full)
    other = mk(3100, 90, 160)                                   # another geometry through the same buffers, then back
    assert np.array_equal(pre.image_adjust_batch(other, (160, 160), mode="letterbox", chunks=3), want(other))
    fr = mk(3200)
    assert np.array_equal(pre.image_adjust_batch(fr, (160, 160), mode="letterbox", chunks=3), want(fr)) and pre.last_d2h_bytes == full
    fa, fb = mk(3300), mk(3400)
    for _ in range(3):
        ha = pre.image_adjust_batch_async(fa, (160, 160), mode="letterbox", chunks=3, copy=False)
        hb = pre.image_adjust_batch_async(fb, (160, 160), mode="letterbox", chunks=3, copy=False)
        assert np.array_equal(ha.result(), want(fa)) and np.array_equal(hb.result(), want(fb))
    assert pre.last_d2h_bytes < full or h == w
    assert not ha.result().flags.writeable


def test_extract_boxes_async_equals_extract_boxes(engine):
    """The pipelined variant: queued on its own stream while a batch preprocess runs on the others; same results, same
    return-the-exception convention."""
    client = Yolov5client(engine=engine)
    pre, post = client.get_preprocess(), client.get_postprocess()
    frames = torch.from_numpy(np.stack([synth.image_noise(180, 320, 700 + i) for i in range(24)])).pin_memory()
    for k in range(3):
        pred = torch.from_numpy(synth.yolo_output(4, 5000, 2, 40 + k)).pin_memory()
        want = post.extract_boxes(HostResponse([pred]), conf_thres=0.3)
        hx = pre.image_adjust_batch_async(frames, (160, 160), mode="letterbox", chunks=4, copy=False)
        pending = post.extract_boxes_async(HostResponse([pred]), conf_thres=0.3)
        x, got = hx.result(), pending.result()
        assert isinstance(got, list) and len(got) == len(want) and pending.result() is got
        for a, b in zip(got, want):
            assert torch.equal(a, b)
        assert np.array_equal(x[3], pre2d.preprocess_frame(frames[3].numpy(), 160, 160, mode="letterbox")[0])
    bg = np.zeros((1, 64, 7), np.float32)
    r = post.extract_boxes_async(HostResponse([bg]), conf_thres=0.3).result()
    assert isinstance(r, AssertionError) and str(r) == str(golden("yolo_post")["empty_msg"])
    r = post.extract_boxes_async(HostResponse([bg]), conf_thres=0.3, iou_thres=2.0).result()
    assert isinstance(r, AssertionError)
    assert isinstance(post.extract_boxes_async(object()).result(), Exception)


def test_two_batches_in_flight(engine):
    """Deferred calls alternate between two sets of host buffers: the second batch is queued before the first result is read."""
    client = Yolov5client(engine=engine)
    pre, post = client.get_preprocess(), client.get_postprocess()
    fa = torch.from_numpy(np.stack([synth.image_noise(180, 320, 800 + i) for i in range(16)])).pin_memory()
    fb = torch.from_numpy(np.stack([synth.image_noise(180, 320, 850 + i) for i in range(16)])).pin_memory()
    pa, pb = (torch.from_numpy(synth.yolo_output(3, 4000, 2, 60 + k)).pin_memory() for k in range(2))
    want = [post.extract_boxes(HostResponse([p]), conf_thres=0.3) for p in (pa, pb)]
    for _ in range(3):
        h1 = (pre.image_adjust_batch_async(fa, (160, 160), mode="letterbox", chunks=4, copy=False), post.extract_boxes_async(HostResponse([pa]), conf_thres=0.3))
        h2 = (pre.image_adjust_batch_async(fb, (160, 160), mode="letterbox", chunks=4, copy=False), post.extract_boxes_async(HostResponse([pb]), conf_thres=0.3))
        xa, da = h1[0].result(), h1[1].result()
        xb, db = h2[0].result(), h2[1].result()
        for frames, x in ((fa, xa), (fb, xb)):
            assert np.array_equal(x, np.concatenate([pre2d.preprocess_frame(f.numpy(), 160, 160, mode="letterbox") for f in frames]))
        for got, w in ((da, want[0]), (db, want[1])):
            assert all(torch.equal(a, b) for a, b in zip(got, w))


def test_two_pageable_batches_in_flight(engine):
    """Deferred calls with ordinary (pageable) numpy frames: each call stages its frames in pinned memory, and the second
    call's staging must not overwrite what the first call's uploads are still reading."""
    pre = Yolov5client(engine=engine).get_preprocess()
    fa = np.stack([synth.image_noise(360, 640, 1800 + i) for i in range(24)])
    fb = np.stack([synth.image_noise(360, 640, 1850 + i) for i in range(24)])
    wa, wb = (np.concatenate([pre2d.preprocess_frame(f, 320, 320, mode="letterbox") for f in fr]) for fr in (fa, fb))
    for _ in range(4):
        h1 = pre.image_adjust_batch_async(fa, (320, 320), mode="letterbox", chunks=6)
        h2 = pre.image_adjust_batch_async(fb, (320, 320), mode="letterbox", chunks=6)
        assert np.array_equal(h1.result(), wa) and np.array_equal(h2.result(), wb)


def test_cuda_graph_replay_matches_eager(engine):
    """The preprocess + YOLO post entry points captured once in a CUDA graph (their launches carry the programmatic
    serialization attribute) and replayed on new inputs give what the eager calls give."""
    dev = engine.device
    frames = torch.zeros((4, 180, 320, 3), dtype=torch.uint8, device=dev)
    pred = torch.zeros((4, 3000, 7), dtype=torch.float32, device=dev)
    x = torch.empty((4, 3, 160, 160), dtype=torch.float32, device=dev)
    det = torch.zeros((4, 300, 6), dtype=torch.float64, device=dev)
    cnt = torch.zeros((2, 4), dtype=torch.int32, device=dev)

    def both():
        engine.preprocess_images(frames, 160, 160, mode="letterbox", div=255.0, out=x)
        engine.yolo_nms(pred, conf_thres=0.3, iou_thres=0.45, max_det=300, out=det, counts=cnt)

    data = [(torch.from_numpy(np.stack([synth.image_noise(180, 320, 300 + 10 * k + i) for i in range(4)])).to(dev),
             torch.from_numpy(synth.yolo_output(4, 3000, 2, 70 + k)).to(dev)) for k in range(3)]
    frames.copy_(data[0][0]); pred.copy_(data[0][1])
    s = torch.cuda.Stream(dev)
    s.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(s):
        both()
    torch.cuda.current_stream(dev).wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        both()
    for f, p in data[1:] + data[:1]:
        frames.copy_(f); pred.copy_(p)
        g.replay()
        torch.cuda.synchronize()
        got = (x.clone(), det.clone(), cnt.clone())
        both()
        torch.cuda.synchronize()
        assert torch.equal(got[0], x) and torch.equal(got[2], cnt)
        for b in range(4):
            n = int(cnt[0, b])
            assert n > 0 and torch.equal(got[1][b, :n], det[b, :n])
