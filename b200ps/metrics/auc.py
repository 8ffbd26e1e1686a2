"""Streaming ROC-AUC with fixed thresholds — the metric lr2.py evaluates with
``tf.contrib.metrics.streaming_auc(predict_op, y)`` (lr2.py:400,455-468).

Semantics kept from TF 0.12 [TF-lib]: ``num_thresholds`` (200) thresholds
``[-1e-7, 1/(T-1), ..., (T-2)/(T-1), 1+1e-7]``; a prediction is positive at threshold t when
``pred > t``; four accumulators TP/FN/TN/FP of shape ``[T]`` persist across ``update`` calls;
the value is the trapezoidal area under (FPR, TPR) with TF's epsilons.

B200 design: one atomic histogram pass on the device (K10, ``auc_hist``) bins each
prediction by the number of thresholds strictly below it; the accumulators at every
threshold are suffix sums of the two histograms, taken on the host when the value is read.
"""
from __future__ import annotations

import numpy as np

_KEPS = 1e-7
_EPS = 1.0e-6


def make_thresholds(num_thresholds: int = 200) -> np.ndarray:
    t = [(i + 1) * 1.0 / (num_thresholds - 1) for i in range(num_thresholds - 2)]
    return np.asarray([0.0 - _KEPS] + t + [1.0 + _KEPS], dtype=np.float32)


def auc_from_counts(tp, fn, tn, fp) -> float:
    tp, fn, tn, fp = (np.asarray(a, dtype=np.float64) for a in (tp, fn, tn, fp))
    rec = (tp + _EPS) / (tp + fn + _EPS)
    fpr = fp / (fp + tn + _EPS)
    n = len(tp)
    return float(np.sum((fpr[: n - 1] - fpr[1:]) * (rec[: n - 1] + rec[1:]) / 2.0))


def auc_from_counts_reference(pred: np.ndarray, labels: np.ndarray, num_thresholds: int = 200):
    """O(B*T) restatement used as the test oracle: explicit ``pred > t`` comparisons."""
    thr = make_thresholds(num_thresholds)
    pred = np.asarray(pred, dtype=np.float32).reshape(-1)
    pos = np.asarray(labels).reshape(-1) > 0.5
    above = pred[None, :] > thr[:, None]
    tp = (above & pos[None, :]).sum(1)
    fp = (above & ~pos[None, :]).sum(1)
    fn = (~above & pos[None, :]).sum(1)
    tn = (~above & ~pos[None, :]).sum(1)
    return tp, fn, tn, fp


class StreamingAUC:
    """``update(pred, labels)`` accumulates; ``result()`` returns the AUC so far.

    ``update`` returns ``(value, update_marker)`` like the ``(value, update_op)`` tuple the
    reference logs raw (lr2.py:461,468).
    """

    def __init__(self, num_thresholds: int = 200, device="cpu"):
        import torch
        self.T = num_thresholds
        self.device = torch.device(device)
        self._thr_np = make_thresholds(num_thresholds)
        self._thr = torch.from_numpy(self._thr_np).to(self.device)
        self.reset()

    def reset(self):
        import torch
        self.hist_pos = torch.zeros(self.T + 1, dtype=torch.int64, device=self.device)
        self.hist_neg = torch.zeros(self.T + 1, dtype=torch.int64, device=self.device)

    def update(self, pred, labels):
        import torch
        pred = pred.reshape(-1).float().contiguous()
        labels = labels.reshape(-1).float().contiguous()
        if pred.is_cuda:
            from .. import ops
            ops.auc_hist(pred, labels, self._thr, self.hist_pos, self.hist_neg)
        else:
            # bucket = number of thresholds strictly below the prediction
            idx = torch.searchsorted(self._thr, pred, right=False)
            pos = labels > 0.5
            self.hist_pos += torch.bincount(idx[pos], minlength=self.T + 1)
            self.hist_neg += torch.bincount(idx[~pos], minlength=self.T + 1)
        return self

    def counts(self):
        hp = self.hist_pos.cpu().numpy().astype(np.int64)
        hn = self.hist_neg.cpu().numpy().astype(np.int64)
        # pred > thr[j]  <=>  bucket > j
        suf_p = np.cumsum(hp[::-1])[::-1]
        suf_n = np.cumsum(hn[::-1])[::-1]
        tp = suf_p[1:]
        fp = suf_n[1:]
        fn = hp.sum() - tp
        tn = hn.sum() - fp
        return tp, fn, tn, fp

    def result(self) -> float:
        return auc_from_counts(*self.counts())
