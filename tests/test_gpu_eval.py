"""GPU parity of the evaluation math (§8f rank 4) through the C ABI: tcb_calculate_metrics / tcb_eval_match /
tcb_ap_per_class against the golden outputs of the reference's own EvaluateInference methods and against the oracle."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import eval_metrics as em
from triton_client_b200 import synth
from triton_client_b200.utils.evaluate_metrics import EvaluationMetrics

pytestmark = pytest.mark.gpu
NFRAMES = 7


def _check(m, ref, k):
    assert np.array_equal(m["ap_class"], ref["ap_class"]), k
    for name in ("p", "r", "f1", "ap"):        # float64, the reference's operation order: identical, not merely close
        assert np.array_equal(m[name], ref[name]), (k, name, np.abs(m[name] - ref[name]).max())


def test_calculate_metrics_golden(engine):
    g = golden("eval")
    ev = EvaluationMetrics(engine)
    seen = {n: [] for n in ("p", "r", "f1", "ap_class")}
    ev.observers = {n: seen[n].append for n in seen}
    ev.all_groundtruths = [{"header": k, "ground_truths": g[f"gts{k}"]} for k in range(NFRAMES)]
    ev.all_predictions = [{"header": k, "prediction": g[f"preds{k}"]} for k in range(NFRAMES)]
    stats = ev.calculate_metrics()
    si = 0
    for k in range(NFRAMES):
        m = ev.frame_metrics[k]
        assert (m is not None) == bool(g[f"matched{k}"])
        if m is None:
            continue
        correct, conf, pcls, tcls = stats[si][0]
        si += 1
        assert np.array_equal(correct.numpy(), g[f"correct{k}"])
        assert np.array_equal(conf, g[f"preds{k}"][:, 4]) and np.array_equal(tcls, g[f"gts{k}"][:, 5])
        _check(m, dict(p=g[f"p{k}"], r=g[f"r{k}"], f1=g[f"f1{k}"], ap=g[f"ap{k}"], ap_class=g[f"cls{k}"]), k)
    assert si == len(stats) == 6
    assert len(seen["p"]) == sum(len(g[f"cls{k}"]) for k in range(NFRAMES) if g[f"matched{k}"])


def test_ap_per_class_golden(engine):
    g = golden("eval")
    ev = EvaluationMetrics(engine)
    for k in (0, 2, 4, 5):
        p, r, ap, f1, cls = ev.ap_per_class(g[f"correct{k}"], g[f"preds{k}"][:, 4], g[f"preds{k}"][:, 5], g[f"gts{k}"][:, 5])
        _check(dict(p=p, r=r, ap=ap, f1=f1, ap_class=cls), dict(p=g[f"p{k}"], r=g[f"r{k}"], f1=g[f"f1{k}"], ap=g[f"ap{k}"], ap_class=g[f"cls{k}"]), k)


@pytest.mark.parametrize("n_gt,nc,frames", [(150, 6, 24), (900, 3, 3), (2, 2, 40)])
def test_calculate_metrics_vs_oracle(engine, n_gt, nc, frames):
    data = [synth.eval_frame(synth.seed_for(10, 100 * n_gt + k), n_gt=n_gt, nc=nc, dup=0.4, false_pos=n_gt // 3 + 1) for k in range(frames)]
    data.append((np.zeros((0, 6)), data[0][1]))             # a frame without ground truths
    data.append((data[0][0], np.zeros((0, 6))))             # ... and one without predictions
    ev = EvaluationMetrics(engine)
    ev.all_groundtruths = [{"ground_truths": d[0]} for d in data]
    ev.all_predictions = [{"prediction": d[1]} for d in data]
    ev.calculate_metrics()
    ref = em.calculate_metrics([d[0] for d in data], [d[1] for d in data])
    nmatched = 0
    for k, (m, r) in enumerate(zip(ev.frame_metrics, ref)):
        assert (m is None) == (r is None), k
        if r is not None:
            nmatched += 1
            _check(m, r, k)
    assert nmatched >= frames // 2


def test_equal_confidences_and_iou_ties(engine):
    """Duplicated prediction rows and labels (equal IoUs, equal confidences): the orders the reference leaves to unstable
    sorts follow the rule the oracle documents (later label on equal IoU, input order on equal confidence)."""
    gts, preds = synth.eval_frame(synth.seed_for(10, 7), n_gt=10, nc=2)
    preds = np.concatenate([preds[:6], preds[:6], preds[6:]], 0)
    gts = np.concatenate([gts, gts[:3]], 0)                 # duplicated labels: the later copy takes the detection
    ev = EvaluationMetrics(engine)
    ev.all_groundtruths, ev.all_predictions = [{"ground_truths": gts}], [{"prediction": preds}]
    stats = ev.calculate_metrics()
    ref = em.calculate_metrics([gts], [preds])[0]
    assert np.array_equal(stats[0][0][0].numpy(), ref["correct"])
    _check(ev.frame_metrics[0], ref, 0)


def test_limits_are_errors(engine):
    from triton_client_b200._lib import TcbError
    big = np.zeros((2049, 6))
    with pytest.raises(TcbError):
        engine.calculate_metrics(torch.from_numpy(big).cuda(), [2049], torch.from_numpy(big).cuda(), [2049])


def test_ap_per_class_edge_frames(engine):
    """No predictions, no targets, a class without predictions, fewer than ten thresholds: same numbers as the oracle."""
    ev = EvaluationMetrics(engine)
    rng = np.random.default_rng(5)
    cases = [(np.zeros((0, 10), bool), np.zeros(0), np.zeros(0), np.array([0.0, 1.0, 1.0])),          # nothing predicted
             (rng.uniform(0, 1, (7, 10)) < 0.5, rng.uniform(0.3, 1, 7), np.zeros(7), np.array([0.0] * 8 + [2.0, 2.0, 5.0])),   # classes 2, 5 unpredicted
             (rng.uniform(0, 1, (40, 3)) < 0.2, rng.uniform(0.3, 1, 40), rng.integers(0, 3, 40).astype(float), rng.integers(0, 3, 60).astype(float)),
             (rng.uniform(0, 1, (1, 10)) < 2.0, np.array([0.9]), np.array([1.0]), np.array([1.0]))]
    for k, (tp, conf, pcls, tcls) in enumerate(cases):
        p, r, ap, f1, cls = ev.ap_per_class(tp, conf, pcls, tcls)
        wp, wr, wap, wf1, wcls = em.ap_per_class(tp, conf, pcls, tcls)
        _check(dict(p=p, r=r, ap=ap, f1=f1, ap_class=cls), dict(p=wp, r=wr, ap=wap, f1=wf1, ap_class=wcls), k)
    out = engine.ap_per_class(torch.zeros((3, 10), dtype=torch.uint8).cuda(), torch.rand(3, dtype=torch.float64).cuda(),
                              torch.zeros(3, dtype=torch.float64).cuda(), [3], torch.zeros(0, dtype=torch.float64).cuda(), [0], max_classes=4)
    assert int(out["num_classes"][0]) == 0                                                              # no targets: no classes
