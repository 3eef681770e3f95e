"""Mirror of the evaluation math of the reference's EvaluateInference (communicator/evaluate_inference.py): same method
names, arguments and return values, host arrays in and out, the arithmetic on the GPU (tcb_calculate_metrics /
tcb_ap_per_class, csrc/eval_metrics.cu). The ROS / Prometheus side of that class is out of scope; `observe`
callbacks stand where its Summary objects are."""
import numpy as np
import torch

from ..clients import wire


class EvaluationMetrics:
    """calculate_metrics / ap_per_class of EvaluateInference (:400-446, :158-218); compute_ap (:131-156) has no method of
    its own here, it runs inside the ap_per_class kernel. All frames of a run go to the device in one upload and are
    evaluated by two kernel launches."""

    def __init__(self, engine=None, observers=None):
        self.engine = engine or wire.default_engine()
        self.all_groundtruths, self.all_predictions = [], []
        self.img_processed = self.gt_processed = True
        self.observers = observers or {}      # name in ('p', 'r', 'f1', 'ap_class') -> callable, like the Summary.observe calls (:440-444)

    def _run(self, gts, preds, max_classes=None):
        eng = self.engine
        gts = [np.asarray(g, np.float64).reshape(-1, 6) for g in gts]
        preds = [np.asarray(p, np.float64).reshape(-1, 6) for p in preds]
        if max_classes is None:
            max_classes = max([1] + [len(np.unique(g[:, 5])) for g in gts])
        gd = torch.from_numpy(np.concatenate(gts + [np.zeros((0, 6))], 0)).to(eng.device)
        pd = torch.from_numpy(np.concatenate(preds + [np.zeros((0, 6))], 0)).to(eng.device)
        out = eng.calculate_metrics(gd, [len(g) for g in gts], pd, [len(p) for p in preds], max_classes)
        return {k: v.cpu().numpy() for k, v in out.items()}

    def calculate_metrics(self):
        """-> aggregated_stats as the reference returns it: per matched frame [(correct, conf, pred_cls, target_cls)];
        per-class p, r, f1 and the class ids go to the observers. Frames without any match are skipped."""
        n = len(self.all_groundtruths)
        gts = [self.all_groundtruths[k]["ground_truths"] for k in range(n)]
        preds = [np.asarray(self.all_predictions[k]["prediction"], np.float64) for k in range(n)]
        o = self._run(gts, preds)
        aggregated, row = [], 0
        self.frame_metrics = []
        for k in range(n):
            d = len(preds[k])
            correct = torch.from_numpy(o["correct"][row:row + d].astype(bool))
            row += d
            if not o["matched"][k]:
                self.frame_metrics.append(None)
                continue
            aggregated.append([(correct, preds[k][:, 4], preds[k][:, 5], np.asarray(gts[k])[:, 5])])
            nc = int(o["num_classes"][k])
            m = dict(p=o["p"][k, :nc], r=o["r"][k, :nc], ap=o["ap"][k, :nc], f1=o["f1"][k, :nc], ap_class=o["classes"][k, :nc])
            self.frame_metrics.append(m)
            for name in ("p", "r", "f1", "ap_class"):
                if name in self.observers:
                    for v in m[name]:
                        self.observers[name](v)
        return aggregated

    def ap_per_class(self, tp, conf, pred_cls, target_cls, plot=False, save_dir='.', names=()):
        eng = self.engine
        tp = np.asarray(tp)
        tp = tp.reshape(len(tp), tp.shape[1] if tp.ndim > 1 else 1)
        target_cls = np.asarray(target_cls, np.float64)
        dev = lambda a, dt: torch.from_numpy(np.ascontiguousarray(np.asarray(a).astype(dt))).to(eng.device)
        nc = max(1, len(np.unique(target_cls)))
        o = eng.ap_per_class(dev(tp, np.uint8), dev(conf, np.float64), dev(pred_cls, np.float64), [len(tp)],
                             dev(target_cls, np.float64), [len(target_cls)], max_classes=nc)
        o = {k: v.cpu().numpy() for k, v in o.items()}
        n = int(o["num_classes"][0])
        return o["p"][0, :n], o["r"][0, :n], o["ap"][0, :n], o["f1"][0, :n], o["classes"][0, :n].astype("int32")
