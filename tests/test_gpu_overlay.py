"""GPU parity for box rescale + overlay drawing through the C ABI. Pixels must be identical to
the oracle (which is pinned against the unmodified RosInference._callback and against cv2, tests/test_oracle_overlay.py)."""
import zlib

import numpy as np
import pytest
import torch

from conftest import golden
from oracle import overlay
from triton_client_b200 import _lib, synth

pytestmark = pytest.mark.gpu

NAMES = ["crop", "weed"]
GREEN = (0, 255, 0)


def _random_det(rng, n, size, nc, spill=20.0):
    x1 = rng.uniform(-spill, size, n)
    y1 = rng.uniform(-spill, size, n)
    wh = np.exp(rng.normal(np.log(size / 6), 0.8, (n, 2)))
    det = np.stack([x1, y1, x1 + wh[:, 0], y1 + wh[:, 1], rng.uniform(0, 1, n), rng.integers(0, nc, n).astype(np.float64)], 1)
    det[::7, 4] = np.round(det[::7, 4] * 200) / 200      # x.xx5 confidences: the '{:.2f}' rounding cases
    return det


def _gpu_draw(engine, frames, det_list, names, color, input_size, **kw):
    b = len(det_list)
    max_det = max(1, max(len(d) for d in det_list))
    det = np.zeros((b, max_det, 6), np.float64)
    cnt = np.zeros(b, np.int32)
    for i, d in enumerate(det_list):
        det[i, :len(d)] = d
        cnt[i] = len(d)
    f = torch.from_numpy(frames).cuda()
    engine.draw_detections(f, torch.from_numpy(det).cuda(), torch.from_numpy(cnt).cuda(), names, color, input_size, **kw)
    return f.cpu().numpy()


@pytest.mark.parametrize("k", [0, 1])
def test_golden_callback_overlay(engine, k):
    g = golden("overlay")
    h, w, size, seed = (int(v) for v in g[f"cfg{k}"])
    frame = synth.image_noise(h, w, seed)
    assert zlib.crc32(frame.tobytes()) == int(g[f"frame_crc{k}"])
    painted = np.unpackbits(g[f"painted{k}"])[:h * w].reshape(h, w).astype(bool)
    for mode in (1, 2):
        out = _gpu_draw(engine, frame[None], [g[f"det{k}"]], NAMES, GREEN, [size, size], math_mode=mode)[0]
        assert np.array_equal((out != frame).any(-1), painted)
        assert (out[painted] == np.array(GREEN, np.uint8)).all()


@pytest.mark.parametrize("thickness,labels,mode", [(2, True, 1), (2, True, 2), (1, False, 1), (1, True, 2), (2, False, 2)])
def test_random_detections_match_oracle(engine, thickness, labels, mode):
    rng = np.random.default_rng(100 * thickness + 10 * labels + mode)
    h, w, size = 150, 260, 128
    frames = np.stack([synth.image_noise(h, w, 50 + i) for i in range(3)])
    names = ["crop", "weed", "traffic light", "Quxz@gp|", ""]
    dets = [_random_det(rng, n, size, len(names)) for n in (12, 0, 31)]
    got = _gpu_draw(engine, frames, dets, names, (255, 0, 128), [size, size + 17], thickness=thickness, labels=labels, math_mode=mode)
    for i in range(3):
        want = overlay.draw_detections(frames[i].copy(), dets[i], names, (255, 0, 128), [size, size + 17], thickness, labels, f64=(mode == 1))
        assert np.array_equal(got[i], want), f"frame {i}: {(got[i] != want).any(-1).sum()} pixels differ"


def test_painting_order_with_per_class_colours_and_many_detections(engine):
    """More detections than one listing pass holds (256), drawn in order: the last one covering a pixel colours it."""
    rng = np.random.default_rng(7)
    h, w, size = 96, 160, 96
    frame = synth.image_noise(h, w, 3)
    table = [(255, 0, 255), (0, 0, 255), (0, 255, 255), (10, 200, 30)]
    det = _random_det(rng, 700, size, 4, spill=5.0)
    got = _gpu_draw(engine, frame[None], [det], ["a", "b", "c", "d"], table, [size, size], thickness=2, labels=False, math_mode=1,
                    per_class_color=True)[0]
    want = overlay.draw_detections(frame.copy(), det, ["a", "b", "c", "d"], table, [size, size], 2, False, True, per_class_color=True)
    assert np.array_equal(got, want)


def test_label_number_formatting(engine):
    """'{:.2f}'.format(conf): ties and near-ties on the exact binary value, drawn and compared as pixels."""
    confs = [0.125, 0.375, 0.005, 0.995, 1.0, 0.0, 0.675, 0.685, 0.015, 0.994999999999, 0.995000000001, 0.5, 12.345, 0.045]
    h, w = 40 * len(confs) + 20, 200
    det = np.array([[5, 30 + 40 * i, 60, 35 + 40 * i, c, 0] for i, c in enumerate(confs)], np.float64)
    frame = np.zeros((h, w, 3), np.uint8)
    got = _gpu_draw(engine, frame[None], [det], ["x"], (1, 2, 3), [w, h], math_mode=1)[0]     # input size = frame size: scale 1
    want = overlay.draw_detections(frame.copy(), det, ["x"], (1, 2, 3), [w, h], 2, True, True)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_scale_boxes(engine, mode):
    rng = np.random.default_rng(mode)
    det = rng.uniform(-30, 700, (257, 6))
    for (oh, ow, in0, in1, normalized) in [(720, 1280, 640, 640, False), (1080, 1920, 512, 384, False), (333, 517, 640, 640, True)]:
        d = det / 700 if normalized else det
        got = engine.scale_boxes(torch.from_numpy(d).cuda(), (oh, ow), (in0, in1), normalized, mode).cpu().numpy()
        if mode == 0:
            want = overlay.scale_box_array(d, (oh, ow), (in0, in1), normalized)
        else:
            want = d.copy()
            for i in range(len(d)):
                want[i, :4] = [float(v) for v in overlay.scale_box(d[i, :4].astype(np.float32), (oh, ow), (in0, in1), normalized, f64=(mode == 1))]
        assert np.array_equal(got, want)


def test_full_size_batch_properties_and_sampled_parity(engine):
    """C2: 64 frames of 1280x720 with the detections of the YOLO post kernel; frames 0 and 63 against the oracle, the
    whole batch through properties: drawing twice changes nothing more, untouched pixels keep their values."""
    b, h, w = 64, 720, 1280
    pred = torch.from_numpy(synth.yolo_output(b, 25200, 2, 5, objects=12, rows_per_object=20)).cuda()
    det, cnt, _ = engine.yolo_nms(pred, conf_thres=0.3)
    assert int(cnt.min()) >= 5
    one = torch.from_numpy(synth.image_noise(h, w, 9)).cuda()
    frames = one.unsqueeze(0).repeat(b, 1, 1, 1).contiguous()
    engine.draw_detections(frames, det, cnt, NAMES, GREEN, [640, 640], math_mode=1)
    once = frames.clone()
    engine.draw_detections(frames, det, cnt, NAMES, GREEN, [640, 640], math_mode=1)
    assert torch.equal(frames, once)
    changed = (once != one.unsqueeze(0)).any(-1)
    assert (once[changed] == torch.tensor(GREEN, dtype=torch.uint8, device="cuda")).all()
    assert 2000 < int(changed[0].sum()) < h * w // 4
    for i in (0, 63):
        want = overlay.draw_detections(one.cpu().numpy().copy(), det[i, :int(cnt[i])].cpu().numpy(), NAMES, GREEN, [640, 640], 2, True, True)
        assert np.array_equal(once[i].cpu().numpy(), want)


def test_strided_frames_and_dropin_overlay(engine):
    from triton_client_b200.clients.overlay import Overlay
    rng = np.random.default_rng(11)
    big = torch.from_numpy(synth.image_noise(200, 300, 4)).cuda()
    view = big[10:150, 20:280, :]                            # row-strided view of a larger buffer
    det = _random_det(rng, 9, 100, 2)
    want = overlay.draw_detections(view.cpu().numpy().copy(), det, NAMES, GREEN, [100, 100], 2, True, True)
    engine.draw_detections(view, torch.from_numpy(det).cuda().unsqueeze(0), None, NAMES, GREEN, [100, 100], math_mode=1)
    assert np.array_equal(view.cpu().numpy(), want)
    ov = Overlay(NAMES, GREEN, [100, 100], math_mode=1, engine=engine)
    host = synth.image_noise(140, 260, 4)
    out = ov.annotate(host, [torch.from_numpy(det)])
    assert np.array_equal(out, overlay.draw_detections(host.copy(), det, NAMES, GREEN, [100, 100], 2, True, True))
    assert np.array_equal(ov.annotate(host, AssertionError("no detections")), host)
    box = det[0, :4].astype(np.float32)
    assert [float(v) for v in ov.scale_boxes(box, (720, 1280))] == [float(v) for v in overlay.scale_box(box, (720, 1280), [100, 100], f64=True)]
    assert np.array_equal(ov.scale_box_array(det, (720, 1280)), overlay.scale_box_array(det, (720, 1280), [100, 100]))


def test_argument_errors(engine):
    f = torch.zeros((1, 8, 8, 3), dtype=torch.uint8, device="cuda")
    d = torch.zeros((1, 1, 6), dtype=torch.float64, device="cuda")
    with pytest.raises(_lib.TcbError):
        engine.draw_detections(f, d, None, NAMES, GREEN, [8, 8], thickness=3)
    with pytest.raises(_lib.TcbError):
        engine.draw_detections(f, d, None, ["n%d" % i for i in range(200)], GREEN, [8, 8])
    with pytest.raises(TypeError):
        engine.draw_detections(f.float(), d, None, NAMES, GREEN, [8, 8])


def test_plot_boxes_cv2_reference_golden(engine):
    """Frames annotated by the unmodified Yolov5postprocess.plot_boxes_cv2 (golden_plot_boxes): per-class colours, class
    names at scale 1.2 / thickness 1, rectangles thickness 1, boxes and labels cut by every frame edge — through the
    drop-in method and its two free-function twins."""
    from triton_client_b200.clients import Yolov5client
    from triton_client_b200.clients import wire
    g = golden("plot_boxes")
    names = [str(v) for v in g["names"]]
    mode = 1 if int(str(g["numpy_version"]).split(".")[0]) < 2 else 2
    post = Yolov5client(engine=engine).get_postprocess()
    for k in range(2):
        img, boxes, cls = g[f"img{k}"], g[f"boxes{k}"], g[f"cls{k}"]
        det = np.concatenate([boxes[:, :4], boxes[:, 5:6], cls[:, None].astype(np.float32)], 1).astype(np.float64)
        got = _gpu_draw(engine, img[None], [det], names, overlay.plot_colors(len(names)), (1, 1), thickness=1, labels="name", math_mode=mode,
                        normalized=True, per_class_color=True, font_scale=1.2, text_thickness=1, text_dy=0)[0]
        assert np.array_equal(got, g[f"out{k}"]), int((got != g[f"out{k}"]).any(-1).sum())
        if (int(np.__version__.split(".")[0]) < 2) == (mode == 1):     # the drop-in follows the running numpy like the reference does
            rows = [[b[0], b[1], b[2], b[3], b[4], b[5], np.int64(ci)] for b, ci in zip(boxes, cls)]
            before = img.copy()
            out = post.plot_boxes_cv2(img, rows, class_names=names)
            assert np.array_equal(out, g[f"out{k}"]) and np.array_equal(img, before)
            wire._engine = engine
            from triton_client_b200.tools import utils as tu
            from triton_client_b200.utils import postprocess as up
            assert np.array_equal(up.plot_boxes_cv2(img, rows, class_names=names), out) and np.array_equal(tu.plot_boxes_cv2(img, rows, class_names=names), out)
            plain = post.plot_boxes_cv2(img, [r[:4] for r in rows])            # no class names: red rectangles only
            want = overlay.draw_detections(img.copy(), det, names, (255, 0, 0), None, thickness=1, labels=False, f64=(mode == 1), normalized=True)
            assert np.array_equal(plain, want)


@pytest.mark.parametrize("scale,thick,dy", [(0.5, 2, -10), (1.2, 1, 0), (0.8, 3, 5), (2.0, 1, -3), (0.35, 1, 0), (1.0, 2, -20)])
def test_text_at_any_scale_and_thickness_matches_oracle(engine, scale, thick, dy):
    """Stroke-level labels: any font scale / thickness, many of them cut by the frame edge."""
    rng = np.random.default_rng(int(scale * 100) + thick)
    h, w, size = 120, 200, 100
    frames = np.stack([synth.image_noise(h, w, 70 + i) for i in range(2)])
    names = ["crop", "weed", "traffic light", "Quxz@gp|", "~{}[]#%&"]
    dets = [_random_det(rng, n, size, len(names), spill=40.0) for n in (14, 9)]
    kw = dict(font_scale=scale, text_thickness=thick, text_dy=dy)
    got = _gpu_draw(engine, frames, dets, names, (9, 200, 77), [size, size], thickness=1, labels=True, math_mode=2, **kw)
    for i in range(2):
        want = overlay.draw_detections(frames[i].copy(), dets[i], names, (9, 200, 77), [size, size], 1, True, f64=False, **kw)
        assert np.array_equal(got[i], want), f"frame {i}: {(got[i] != want).any(-1).sum()} pixels differ"
