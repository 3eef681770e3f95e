"""Oracle: evaluation math. Test infrastructure only.

Restates, in plain numpy float64 with explicit loops, what communicator/evaluate_inference.py does per frame:
  calculate_metrics :400-446  IoU(ground truths x predictions) -> one-to-one matching -> `correct` (D,10) at the ten
                              IoU thresholds 0.5 .. 0.95, then ap_per_class on that frame's rows
  ap_per_class      :158-218  per target class: cumulative TP/FP in descending confidence, recall/precision sampled
                              on 1000 confidence points, AP per threshold, F1, values at the best mean-F1 point
  compute_ap        :131-156  precision envelope, 101-point interpolation, trapezoid rule
  box_iou           clients/postprocess/base_postprocess.py:48-70 (float64 here: both inputs are float64 arrays)

Pinned by tests/golden/eval.npz: the unmodified EvaluateInference methods run on the same frames
(oracle/make_golden.py: golden_eval).

Where the reference leaves an order to an unstable sort, this file fixes it:
  * matching, `matches[:, 2].argsort()[::-1]` (:419): equal IoUs are taken in descending (label, detection) position,
    which is what a stable ascending sort read backwards gives;
  * `np.argsort(-conf)` (:174): equal confidences keep their original order.
"""
import numpy as np

IOUV = np.array([0.5, 0.55, 0.6, 0.65, 0.70, 0.75, 0.80, 0.85, 0.9, 0.95], np.float32)   # torch.tensor([...]) is float32
NPX = 1000


def box_iou(a, b):
    """(G,4) x (D,4) float64 xyxy -> (G,D); operation order of base_postprocess.py:61-70 (elementwise, so numpy
    broadcasting rounds exactly like a scalar loop)."""
    area1 = (a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1])
    area2 = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
    iw = np.maximum(np.minimum(a[:, None, 2], b[None, :, 2]) - np.maximum(a[:, None, 0], b[None, :, 0]), 0.0)
    ih = np.maximum(np.minimum(a[:, None, 3], b[None, :, 3]) - np.maximum(a[:, None, 1], b[None, :, 1]), 0.0)
    inter = iw * ih
    with np.errstate(invalid="ignore", divide="ignore"):
        return inter / (area1[:, None] + area2[None, :] - inter)


def match_frame(gts, preds):
    """evaluate_inference.py:408-425. Returns (correct (D,10) bool, matched) — `matched` False when no pair passes
    IoU >= 0.5 with equal class (the reference then records nothing for the frame)."""
    g, d = len(gts), len(preds)
    correct = np.zeros((d, len(IOUV)), bool)
    if g == 0 or d == 0:
        return correct, False
    iou = box_iou(gts[:, :4], preds[:, :4])
    ok = (iou >= 0.5) & (gts[:, 5][:, None] == preds[:, 5][None, :])
    best_label = np.full(d, -1)
    for j in np.nonzero(ok.any(0))[0]:      # first np.unique: per detection the highest IoU; ties -> later label
        for i in np.nonzero(ok[:, j])[0]:
            if best_label[j] < 0 or iou[i, j] >= iou[best_label[j], j]:
                best_label[j] = i
    if (best_label < 0).all():
        return correct, False
    taken = set()
    for j in range(d):                      # second np.unique (rows now in detection order): per label the first detection
        i = best_label[j]
        if i < 0 or i in taken:
            continue
        taken.add(i)
        correct[j] = np.float32(iou[i, j]) >= IOUV          # torch.Tensor(matches) is float32
    return correct, True


def _interp(x, xp, fp, left, right):
    """numpy.interp for increasing xp (with repeats): j = last index with xp[j] <= x; slope*(x - xp[j]) + fp[j]."""
    n = len(xp)
    out = np.empty(len(x), np.float64)
    for k, xv in enumerate(x):
        if n == 1:
            out[k] = left if xv < xp[0] else (right if xv > xp[0] else fp[0])
            continue
        if xv < xp[0]:
            out[k] = left
        elif xv > xp[n - 1]:
            out[k] = right
        else:
            j = int(np.searchsorted(xp, xv, side="right")) - 1
            if j == n - 1 or xp[j] == xv:
                out[k] = fp[j]
            else:
                slope = (fp[j + 1] - fp[j]) / (xp[j + 1] - xp[j])
                out[k] = slope * (xv - xp[j]) + fp[j]
    return out


def _pairwise_sum(v):
    """numpy's float64 add.reduce of a contiguous vector shorter than 128: eight running sums, then the rest in order."""
    n = len(v)
    if n < 8:
        s = 0.0
        for t in v:
            s += t
        return s
    r = [v[k] for k in range(8)]
    i = 8
    while i < n - (n % 8):
        for k in range(8):
            r[k] = r[k] + v[i + k]
        i += 8
    res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
    while i < n:
        res += v[i]
        i += 1
    return res


def compute_ap(recall, precision):
    """evaluate_inference.py:131-156 ('interp' method)."""
    mrec = np.concatenate(([0.0], recall, [1.0]))
    mpre = np.concatenate(([1.0], precision, [0.0]))
    for k in range(len(mpre) - 2, -1, -1):
        mpre[k] = max(mpre[k], mpre[k + 1])
    x = np.arange(101) * (1.0 / 100)
    x[-1] = 1.0
    y = _interp(x, mrec, mpre, mpre[0], mpre[-1])
    terms = np.array([(x[k + 1] - x[k]) * (y[k + 1] + y[k]) / 2.0 for k in range(100)])
    return _pairwise_sum(terms)


def ap_per_class(tp, conf, pred_cls, target_cls):
    """evaluate_inference.py:158-218 -> p (nc,), r (nc,), ap (nc,10), f1 (nc,), classes (nc,) int32."""
    order = np.argsort(-conf, kind="stable")
    tp, conf, pred_cls = tp[order], conf[order], pred_cls[order]
    classes = np.unique(target_cls)
    nc, nt = len(classes), tp.shape[1]
    px = np.arange(NPX) * (1.0 / (NPX - 1))
    px[-1] = 1.0
    ap, p, r = np.zeros((nc, nt)), np.zeros((nc, NPX)), np.zeros((nc, NPX))
    for ci, c in enumerate(classes):
        sel = pred_cls == c
        n_l, n_p = int((target_cls == c).sum()), int(sel.sum())
        if n_p == 0 or n_l == 0:
            continue
        tpc = np.cumsum(tp[sel].astype(np.float64), 0)
        fpc = np.cumsum(1 - tp[sel].astype(np.float64), 0)
        recall = tpc / (n_l + 1e-16)
        precision = tpc / (tpc + fpc)
        xc = -conf[sel]
        r[ci] = _interp(-px, xc, recall[:, 0], 0.0, recall[-1, 0])
        p[ci] = _interp(-px, xc, precision[:, 0], 1.0, precision[-1, 0])
        for j in range(nt):
            ap[ci, j] = compute_ap(recall[:, j], precision[:, j])
    f1 = 2 * p * r / (p + r + 1e-16)
    mean = np.zeros(NPX)
    for ci in range(nc):
        mean = mean + f1[ci]
    i = int(np.argmax(mean / nc)) if nc else 0
    return p[:, i], r[:, i], ap, f1[:, i], classes.astype("int32")


def calculate_metrics(all_gts, all_preds):
    """Per frame (evaluate_inference.py:404-444): None when nothing matched, else dict(correct, p, r, ap, f1, ap_class)."""
    out = []
    for gts, preds in zip(all_gts, all_preds):
        correct, matched = match_frame(gts, preds)
        if not matched:
            out.append(None)
            continue
        p, r, ap, f1, cls = ap_per_class(correct, preds[:, 4], preds[:, 5], gts[:, 5])
        out.append(dict(correct=correct, p=p, r=r, ap=ap, f1=f1, ap_class=cls))
    return out
