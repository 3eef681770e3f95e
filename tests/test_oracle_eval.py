"""Oracle pinning for the evaluation math (§8f rank 4) against golden outputs of the reference's own
EvaluateInference.calculate_metrics / ap_per_class / compute_ap (oracle/make_golden.py: golden_eval)."""
import numpy as np

from conftest import golden
from oracle import eval_metrics as em

NFRAMES = 7


def _frames(g):
    return [(g[f"gts{k}"], g[f"preds{k}"]) for k in range(NFRAMES)]


def test_golden_calculate_metrics():
    g = golden("eval")
    frames = _frames(g)
    res = em.calculate_metrics([f[0] for f in frames], [f[1] for f in frames])
    seen = 0
    for k, r in enumerate(res):
        assert (r is not None) == bool(g[f"matched{k}"])
        if r is None:
            continue
        seen += 1
        assert np.array_equal(r["correct"], g[f"correct{k}"])
        assert np.array_equal(r["ap_class"], g[f"cls{k}"])
        for name in ("p", "r", "f1", "ap"):                      # float64 throughout, same operation order: identical
            assert np.array_equal(r[name], g[f"{name}{k}"]), (k, name)
    assert seen >= 5


def test_interp_matches_numpy():
    rng = np.random.default_rng(3)
    for n in (1, 2, 7, 50):
        xp = np.sort(rng.uniform(-1, 0, n))
        if n > 3:
            xp[2] = xp[1]                                         # repeated abscissa (two rows with equal confidence)
        fp = rng.uniform(0, 1, n)
        x = np.concatenate([rng.uniform(-1.2, 0.2, 200), xp])
        assert np.array_equal(em._interp(x, xp, fp, 0.25, fp[-1]), np.interp(x, xp, fp, left=0.25))


def test_pairwise_sum_matches_numpy():
    rng = np.random.default_rng(4)
    for n in (3, 8, 100, 101, 127):
        v = rng.uniform(0, 1, n)
        assert em._pairwise_sum(v) == v.sum()
