"""GPU parity for the 3D postprocess through the C ABI. Keep lists, labels and counts must be identical;
float boxes/scores within north_star's 1e-5 relative (transcendentals differ by ulps between libm and CUDA)."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import box3d
from triton_client_b200 import synth

pytestmark = pytest.mark.gpu

KITTI_ANCHORS = dict(anchor_range=[0, -39.68, -3, 69.12, 39.68, 1],
                     anchor_sizes=[[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]],
                     anchor_bottoms=[-1.78, -0.6, -0.6], anchor_rotations=[[0, 1.57]] * 3)
NUSC = dict(point_cloud_range=[-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], voxel_size=[0.2, 0.2, 8], feature_map_stride=4,
            post_center_limit_range=[-61.2, -61.2, -10.0, 61.2, 61.2, 10.0])
CLASS_MAP = [[0], [1, 2], [3, 4], [5], [6, 7], [8, 9]]


def _close(a, b, rtol=1e-5, atol=1e-6):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


def test_rotated_iou_matrix(engine):
    a, b = synth.random_boxes7(500, 11, extent=25), synth.random_boxes7(600, 12, extent=25)
    b[:50] = a[:50]
    b[50:100, :2] += 0.02
    want = box3d.boxes_iou_bev(a, b)
    got = engine.rotated_iou(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()).cpu().numpy()
    assert (want > 0).sum() > 5000
    # sin/cos/atan2 differ by ulps; the polygon area is continuous in them
    _close(got, want, rtol=1e-4, atol=2e-6)


# above 512 candidates the mask is built by the whole grid and swept by one CTA; at or below, one CTA does everything
@pytest.mark.parametrize("n,thresh,pre,post,extent", [
    (4096, 0.01, 4096, 500, 60), (3000, 0.2, 1000, 83, 60), (9000, 0.1, 4096, 500, 60), (50, 0.5, 4096, 500, 60), (1, 0.1, 10, 10, 60),
    (513, 0.3, 4096, 500, 20), (512, 0.3, 4096, 500, 20), (1000, 0.05, 777, 1000, 30), (4096, 0.9, 4096, 1024, 80),
    (4096, 0.0, 4096, 500, 12), (2111, 0.1, 4096, 64, 25), (5000, 0.3, 641, 640, 40)])
def test_rotated_nms_keep_lists(engine, n, thresh, pre, post, extent):
    boxes = synth.random_boxes7(n, 20 + n, extent=extent)
    scores = np.round(np.random.default_rng(n).uniform(0, 1, n) * 4096) / 4096    # ties -> index order
    scores = scores.astype(np.float32)
    sel, _ = box3d.class_agnostic_nms(torch.from_numpy(scores), torch.from_numpy(boxes), thresh, pre, post)
    keep, cnt = engine.rotated_nms(torch.from_numpy(boxes).cuda(), torch.from_numpy(scores).cuda(), thresh, pre, post)
    k = int(cnt.item())
    assert keep[:k].cpu().tolist() == sel.tolist()


@pytest.mark.parametrize("n,thresh,pre,post,extent", [(300, 0.1, 4096, 500, 12), (1000, 0.05, 777, 1000, 30), (4096, 0.5, 4096, 1024, 40),
                                                     (4096, 0.0, 4096, 500, 12), (5000, 0.3, 641, 640, 25)])
def test_nms_normal_keep_lists(engine, n, thresh, pre, post, extent):
    """NMS_TYPE nms_normal_gpu (axis-aligned BEV IoU, heading ignored): the CTA path (<= 512 candidates) and the wide one."""
    boxes = synth.random_boxes7(n, 77 + n, extent=extent)
    scores = (np.round(np.random.default_rng(n + 1).uniform(0, 1, n) * 4096) / 4096).astype(np.float32)
    sel, _ = box3d.class_agnostic_nms(torch.from_numpy(scores), torch.from_numpy(boxes), thresh, pre, post, None, "nms_normal_gpu")
    keep, cnt = engine.rotated_nms(torch.from_numpy(boxes).cuda(), torch.from_numpy(scores).cuda(), thresh, pre, post, nms_type="nms_normal_gpu")
    assert keep[:int(cnt.item())].cpu().tolist() == sel.tolist()
    rot, _ = box3d.class_agnostic_nms(torch.from_numpy(scores), torch.from_numpy(boxes), thresh, pre, post)
    assert n < 1000 or rot.tolist() != sel.tolist()        # the two NMS types do differ on these boxes
    with pytest.raises(AttributeError):
        engine.rotated_nms(torch.from_numpy(boxes).cuda(), torch.from_numpy(scores).cuda(), thresh, pre, post, nms_type="nms_cpu")


def test_pp_and_centerpoint_posts_with_nms_normal(engine):
    """NMS_CONFIG.NMS_TYPE = nms_normal_gpu through both head posts (full-size frames)."""
    cls, box, dirp = synth.pp_head_outputs(5201)
    want = box3d.pp_anchor_post(cls, box, dirp, nms_type="nms_normal_gpu", **KITTI_ANCHORS)
    gb, gs, gl = _pp_gpu(engine, cls, box, dirp, nms_type="nms_normal_gpu", **KITTI_ANCHORS)
    assert np.array_equal(gl, want["pred_labels"].numpy()) and len(gl) > 20
    _close(gs, want["pred_scores"].numpy())
    _close(gb, want["pred_boxes"].numpy(), rtol=1e-5, atol=1e-5)
    rot = box3d.pp_anchor_post(cls, box, dirp, **KITTI_ANCHORS)
    assert len(rot["pred_labels"]) != len(gl) or not np.array_equal(rot["pred_scores"].numpy(), want["pred_scores"].numpy())
    heads = synth.centerpoint_head_outputs(5202)
    want = box3d.centerhead_generate_predicted_boxes(heads, CLASS_MAP, nms_type="nms_normal_gpu", **NUSC)
    gb, gs, gl = _cp_gpu(engine, heads, nms_type="nms_normal_gpu")
    assert np.array_equal(gl, want["pred_labels"].numpy()) and len(gl) > 20
    _close(gs, want["pred_scores"].numpy())


def test_rotated_nms_empty(engine):
    keep, cnt = engine.rotated_nms(torch.zeros((0, 7)).cuda(), torch.zeros((0,)).cuda(), 0.1)
    assert int(cnt.item()) == 0


def _pp_gpu(engine, cls, box, dirp, **kw):
    ob, osc, ol, cnt = engine.pp_anchor_post(torch.from_numpy(cls).cuda(), torch.from_numpy(box).cuda(),
                                             torch.from_numpy(dirp).cuda() if dirp is not None else None, **kw)
    k = int(cnt.item())
    return ob[:k].cpu().numpy(), osc[:k].cpu().numpy(), ol[:k].cpu().numpy()


def test_pp_golden_decode_pieces(engine):
    """Anchors + ResidualCoder decode + direction bins against the reference's golden vectors: feed encodings for a
    handful of anchors with huge scores so that every one of them comes back, then compare box by box."""
    g = golden("box3d")
    h, w = 248, 216
    rng = np.random.default_rng(3)
    cls = np.full((h, w, 18), -8.0, np.float32)
    box = rng.normal(0, 0.3, (h, w, 42)).astype(np.float32)
    idx = g["anchor_idx"][::40][:90]
    flat_cls = cls.reshape(-1, 3)
    flat_cls[idx, 0] = 8.0 - np.arange(len(idx)) / 64.0   # distinct, descending scores
    far = box3d.generate_anchors((w, h), **{k: KITTI_ANCHORS[k] for k in ("anchor_range", "anchor_sizes", "anchor_rotations", "anchor_bottoms")})
    gb, gs, gl = _pp_gpu(engine, cls, box, None, score_thresh=0.5, nms_thresh=1.0, **KITTI_ANCHORS)   # thresh 1.0: nothing suppressed
    assert len(gs) == len(idx) and (gl == 1).all()
    want = box3d.decode_torch(torch.from_numpy(box.reshape(-1, 7)[idx]), far[idx]).numpy()
    _close(far[g["anchor_idx"]].numpy(), g["anchor_val"], rtol=1e-6, atol=1e-6)
    _close(gb, want, rtol=1e-5, atol=1e-6)
    assert np.array_equal(gb[:, [2, 6]], want[:, [2, 6]])   # no transcendental, no arange on these: bit-exact


@pytest.mark.parametrize("seed", [4001, 4002])
def test_pp_full_frame_c4(engine, seed):
    cls, box, dirp = synth.pp_head_outputs(seed)
    want = box3d.pp_anchor_post(cls, box, dirp, **KITTI_ANCHORS)
    gb, gs, gl = _pp_gpu(engine, cls, box, dirp, **KITTI_ANCHORS)
    assert gb.shape == tuple(want["pred_boxes"].shape)
    assert np.array_equal(gl, want["pred_labels"].numpy())
    _close(gs, want["pred_scores"].numpy())
    _close(gb, want["pred_boxes"].numpy(), rtol=1e-5, atol=1e-5)


def test_pp_no_candidates_and_no_dir(engine):
    cls = np.full((20, 24, 18), -8.0, np.float32)
    box = np.zeros((20, 24, 42), np.float32)
    gb, gs, gl = _pp_gpu(engine, cls, box, None, **KITTI_ANCHORS)
    assert gb.shape == (0, 7)
    cls2, box2, _ = synth.pp_head_outputs(5, h=20, w=24, planted=6)
    want = box3d.pp_anchor_post(cls2, box2, None, **KITTI_ANCHORS)
    gb, gs, gl = _pp_gpu(engine, cls2, box2, None, **KITTI_ANCHORS)
    assert np.array_equal(gl, want["pred_labels"].numpy())
    _close(gb, want["pred_boxes"].numpy(), rtol=1e-5, atol=1e-5)


def _cp_gpu(engine, heads, class_map=None, **kw):
    dev = [{k: torch.from_numpy(v).cuda() for k, v in hd.items()} for hd in heads]
    ob, osc, ol, cnt = engine.centerpoint_post(dev, class_map or CLASS_MAP[:len(heads)], NUSC["feature_map_stride"], NUSC["voxel_size"],
                                               NUSC["point_cloud_range"], NUSC["post_center_limit_range"], **kw)
    k = int(cnt.item())
    return ob[:k].cpu().numpy(), osc[:k].cpu().numpy(), ol[:k].cpu().numpy()


@pytest.mark.parametrize("seed", [5001, 5002])
def test_centerpoint_full_frame_c5(engine, seed):
    heads = synth.centerpoint_head_outputs(seed)
    want = box3d.centerhead_generate_predicted_boxes(heads, CLASS_MAP, **NUSC)
    gb, gs, gl = _cp_gpu(engine, heads)
    assert gb.shape == tuple(want["pred_boxes"].shape)
    assert np.array_equal(gl, want["pred_labels"].numpy())
    _close(gs, want["pred_scores"].numpy())
    _close(gb, want["pred_boxes"].numpy(), rtol=1e-5, atol=1e-5)


def test_centerpoint_golden_decode(engine):
    """decode_bbox_from_heatmap golden (reference code, tie-free heatmap): NMS disabled (thresh 1.0), K=40."""
    g = golden("box3d")
    p = 1e-7
    hm = np.log(np.clip(g["cp_hm"][0], p, 1 - p) / (1 - np.clip(g["cp_hm"][0], p, 1 - p))).astype(np.float32)
    head = {"hm": hm, "center": g["cp_center"][0], "center_z": g["cp_center_z"][0], "dim": np.log(g["cp_dim"][0]).astype(np.float32),
            "rot": np.concatenate([g["cp_rot_cos"][0], g["cp_rot_sin"][0]]), "vel": g["cp_vel"][0]}
    gb, gs, gl = _cp_gpu(engine, [head], class_map=[[0, 1]], k=40, nms_thresh=2.0, pre_max=1000, post_max=83)  # 2.0: identical boxes of two classes can reach IoU 1+ulp
    # logit(sigmoid(x)) is not exact, so compare as sets of decoded rows ordered by score with a loose score tolerance
    assert gb.shape == g["cp_boxes"].shape
    _close(gb[:, [0, 1, 2, 6, 7, 8]], g["cp_boxes"][:, [0, 1, 2, 6, 7, 8]], rtol=1e-5, atol=1e-5)
    _close(gb[:, 3:6], g["cp_boxes"][:, 3:6], rtol=1e-4)
    _close(gs, g["cp_scores"], rtol=1e-4)
    assert np.array_equal(gl - 1, g["cp_labels"])


def test_class_score_filter(engine):
    rng = np.random.default_rng(1)
    labels = rng.integers(0, 10, 400).astype(np.int64)
    scores = (np.round(rng.uniform(0, 1, 400) * 20) / 20).astype(np.float32)
    want = box3d.remove_low_score_nu({"labels": labels, "scores": scores, "i": np.arange(400)})["i"]
    idx, cnt = engine.class_score_filter(torch.from_numpy(labels).cuda(), torch.from_numpy(scores).cuda(), box3d.NUSC_GROUPS)
    assert idx[:int(cnt.item())].cpu().tolist() == want.tolist()


def test_class_score_filter_tables_queued_back_to_back(engine):
    """The group tables travel through the context's pinned slots (eight of them, reused in turn): twenty calls with
    twenty different tables queued without a sync in between must each see their own table."""
    rng = np.random.default_rng(2)
    labels = torch.from_numpy(rng.integers(0, 10, 300).astype(np.int64)).cuda()
    scores_h = (np.round(rng.uniform(0, 1, 300) * 50) / 50).astype(np.float32)
    scores = torch.from_numpy(scores_h).cuda()
    tables = [[(int(c), float(np.float32(0.02 * (k + c)))) for c in rng.permutation(10)[:1 + k % 7]] for k in range(20)]
    res = [engine.class_score_filter(labels, scores, t) for t in tables]
    torch.cuda.synchronize()
    lab = labels.cpu().numpy()
    for t, (idx, cnt) in zip(tables, res):
        want = [i for c, th in t for i in np.nonzero((lab == c) & (scores_h >= np.float32(th)))[0].tolist()]
        assert idx[:int(cnt.item())].cpu().tolist() == want


PP_SMALL = dict(anchor_sizes=KITTI_ANCHORS["anchor_sizes"], anchor_bottoms=KITTI_ANCHORS["anchor_bottoms"],
                anchor_rotations=KITTI_ANCHORS["anchor_rotations"])


def _same_as_reference(gb, gs, gl, g, tag, b):
    """Against outputs of the unmodified pcdet code: rows, order and labels identical, floats within 1e-5."""
    assert gb.shape == g[f"{tag}_boxes{b}"].shape
    assert np.array_equal(gl, g[f"{tag}_labels{b}"])
    _close(gs, g[f"{tag}_scores{b}"])
    _close(gb, g[f"{tag}_boxes{b}"], rtol=1e-5, atol=1e-5)


def test_pp_reference_glue_golden(engine):
    """The whole PointPillars chain as pcdet runs it (AnchorHeadTemplate.generate_predicted_boxes -> post_processing ->
    class_agnostic_nms -> nms_gpu, unmodified: oracle/make_golden.py golden_pcdet_post), batch of two."""
    g = golden("pcdet_post")
    for b in range(2):
        gb, gs, gl = _pp_gpu(engine, g["pp_cls"][b], g["pp_box"][b], g["pp_dir"][b], anchor_range=g["pp_range"].tolist(), **PP_SMALL)
        _same_as_reference(gb, gs, gl, g, "pp", b)
    # the batch in one call, as generate_predicted_boxes(batch_size, ...) / post_processing take it
    ob, osc, ol, cnt = engine.pp_anchor_post(*(torch.from_numpy(g[k]).cuda() for k in ("pp_cls", "pp_box", "pp_dir")),
                                             anchor_range=g["pp_range"].tolist(), **PP_SMALL)
    assert ob.shape == (2, 500, 7) and cnt.shape == (2,)
    for b in range(2):
        k = int(cnt[b])
        _same_as_reference(ob[b, :k].cpu().numpy(), osc[b, :k].cpu().numpy(), ol[b, :k].cpu().numpy(), g, "pp", b)


def test_centerpoint_reference_glue_golden(engine):
    """CenterHead.generate_predicted_boxes run unmodified on a batch of two, two heads (1 + 2 classes), K=60, post 30."""
    g = golden("pcdet_post")
    for b in range(2):
        heads = [{k: np.ascontiguousarray(g[f"cp{h}_{k}"][b]) for k in ("hm", "center", "center_z", "dim", "rot", "vel")} for h in range(2)]
        gb, gs, gl = _cp_gpu(engine, heads, class_map=[[0], [1, 2]], k=60, score_thresh=0.1, nms_thresh=0.2, pre_max=1000, post_max=30)
        _same_as_reference(gb, gs, gl, g, "cp", b)
    heads = [{k: torch.from_numpy(g[f"cp{h}_{k}"]).cuda() for k in ("hm", "center", "center_z", "dim", "rot", "vel")} for h in range(2)]
    ob, osc, ol, cnt = engine.centerpoint_post(heads, [[0], [1, 2]], NUSC["feature_map_stride"], NUSC["voxel_size"], NUSC["point_cloud_range"],
                                               NUSC["post_center_limit_range"], k=60, score_thresh=0.1, nms_thresh=0.2, pre_max=1000, post_max=30)
    assert ob.shape == (2, 60, 9) and cnt.shape == (2,)
    for b in range(2):
        k = int(cnt[b])
        _same_as_reference(ob[b, :k].cpu().numpy(), osc[b, :k].cpu().numpy(), ol[b, :k].cpu().numpy(), g, "cp", b)


@pytest.mark.parametrize("tag,nms_type", [("rot", "nms_gpu"), ("rot_all", "nms_gpu"), ("normal", "nms_normal_gpu")])
def test_class_agnostic_nms_reference_golden(engine, tag, nms_type):
    """model_nms_utils.class_agnostic_nms run unmodified (score mask -> topk -> NMS_TYPE -> post cut -> index map)."""
    g = golden("pcdet_post")
    thr, st, pre, post = g[f"nms_{tag}_cfg"]
    boxes, scores = g["nms_boxes"], g["nms_scores"]
    idx = np.nonzero(scores >= np.float32(st))[0] if st >= 0 else np.arange(len(scores))
    keep, cnt = engine.rotated_nms(torch.from_numpy(boxes[idx]).cuda(), torch.from_numpy(scores[idx]).cuda(), float(thr), int(pre), int(post),
                                   nms_type=nms_type)
    sel = idx[keep[:int(cnt.item())].cpu().numpy()]
    assert np.array_equal(sel, g[f"nms_{tag}_sel"])


@pytest.mark.parametrize("tag,nms_type", [("rot", "nms_gpu"), ("normal", "nms_normal_gpu")])
def test_multi_classes_nms_reference_golden(engine, tag, nms_type):
    """model_nms_utils.multi_classes_nms run unmodified: three classes, rows with two extra columns riding along."""
    from triton_client_b200.clients import PointPillarPostprocess
    g = golden("pcdet_post")
    thr, st, pre, post = g[f"mc_{tag}_cfg"]
    ps, pl, pb = engine.multi_classes_nms(torch.from_numpy(g["mc_scores"]).cuda(), torch.from_numpy(g["mc_boxes"]).cuda(), float(thr), int(pre), int(post),
                                          None if st < 0 else float(st), nms_type)
    assert np.array_equal(ps.cpu().numpy(), g[f"mc_{tag}_scores"]) and np.array_equal(pl.cpu().numpy(), g[f"mc_{tag}_labels"])
    assert np.array_equal(pb.cpu().numpy(), g[f"mc_{tag}_boxes"])
    qs, ql, qb = PointPillarPostprocess(engine=engine).multi_classes_nms(g["mc_scores"], g["mc_boxes"], float(thr), int(pre), int(post),
                                                                        None if st < 0 else float(st), nms_type)
    assert np.array_equal(qs.numpy(), g[f"mc_{tag}_scores"]) and np.array_equal(ql.numpy(), g[f"mc_{tag}_labels"]) and np.array_equal(qb.numpy(), g[f"mc_{tag}_boxes"])
    # a cut nobody passes, and a class nobody passes
    es, el, eb = engine.multi_classes_nms(torch.from_numpy(g["mc_scores"]).cuda(), torch.from_numpy(g["mc_boxes"]).cuda(), 0.1, 100, 10, 2.0, nms_type)
    assert es.numel() == 0 and el.numel() == 0 and eb.shape == (0, 9)


def test_remove_low_score_nu_reference_golden(engine):
    from triton_client_b200.clients import PointPillarPostprocess
    g = golden("low_score_nu")
    post = PointPillarPostprocess(engine=engine)
    res = post.remove_low_score_nu({"box3d_lidar": g["boxes"], "scores": g["scores"], "labels": g["labels"], "metadata": {}}, 0.45)
    assert "metadata" not in res
    assert np.array_equal(res["box3d_lidar"], g["out_boxes"]) and np.array_equal(res["scores"], g["out_scores"]) and np.array_equal(res["labels"], g["out_labels"])
    assert post.get_annotations_indices(6, 0.15, g["labels"], g["scores"]) == g["motorcycle_idx"].tolist()


def test_batched_frames_equal_their_single_calls(engine):
    """Full-size C4 / C5 frames in one call (wide NMS path, several frames): every frame's rows are those of its own call."""
    frames = [synth.pp_head_outputs(4100 + i) for i in range(3)]
    cls, box, dirp = (torch.from_numpy(np.stack([f[j] for f in frames])).cuda() for j in range(3))
    ob, osc, ol, cnt = engine.pp_anchor_post(cls, box, dirp, **KITTI_ANCHORS)
    for i, f in enumerate(frames):
        gb, gs, gl = _pp_gpu(engine, *f, **KITTI_ANCHORS)
        k = int(cnt[i])
        assert k == len(gs) and k > 20
        assert np.array_equal(ob[i, :k].cpu().numpy(), gb) and np.array_equal(osc[i, :k].cpu().numpy(), gs) and np.array_equal(ol[i, :k].cpu().numpy(), gl)
    want = box3d.pp_anchor_post(*frames[2], **KITTI_ANCHORS)
    k = int(cnt[2])
    assert np.array_equal(ol[2, :k].cpu().numpy(), want["pred_labels"].numpy())
    _close(ob[2, :k].cpu().numpy(), want["pred_boxes"].numpy(), rtol=1e-5, atol=1e-5)
    cps = [synth.centerpoint_head_outputs(5100 + i) for i in range(3)]
    heads = [{k: torch.from_numpy(np.stack([c[h][k] for c in cps])).cuda() for k in cps[0][h]} for h in range(6)]
    ob, osc, ol, cnt = engine.centerpoint_post(heads, CLASS_MAP, NUSC["feature_map_stride"], NUSC["voxel_size"], NUSC["point_cloud_range"],
                                               NUSC["post_center_limit_range"])
    for i, c in enumerate(cps):
        gb, gs, gl = _cp_gpu(engine, c)
        k = int(cnt[i])
        assert k == len(gs) and k > 20
        assert np.array_equal(ob[i, :k].cpu().numpy(), gb) and np.array_equal(osc[i, :k].cpu().numpy(), gs) and np.array_equal(ol[i, :k].cpu().numpy(), gl)
    # rotated NMS of several box sets at once (single-CTA path and wide path)
    for n in (300, 3000):
        boxes = np.stack([synth.random_boxes7(n, 600 + i, extent=30.0) for i in range(4)])
        scores = np.stack([np.random.default_rng(700 + i).permutation(n).astype(np.float32) / n for i in range(4)])
        keep, cnt = engine.rotated_nms(torch.from_numpy(boxes).cuda(), torch.from_numpy(scores).cuda(), 0.1, 4096, 500)
        for i in range(4):
            want = box3d.nms_gpu(boxes[i], scores[i], 0.1, pre_maxsize=4096)[:500].tolist()
            assert keep[i, :int(cnt[i])].cpu().tolist() == want
