"""Host (numpy) evaluation metrics: logistic loss, MSE and tie-aware ROC-AUC.

These restate ``util.py:136-141`` and the ``roc_curve`` + ``auc`` pair used by every
evaluation epilogue (naive.py:194-197).  They are the float64 reference for the device
kernels in ``ops/metrics.py``.
"""
from __future__ import annotations

import numpy as np


def logistic_loss(y, p) -> float:
    """``sum(log(1 + exp(-y p))) / n`` (util.py:136-137), overflow-safe."""
    return float(np.mean(np.logaddexp(0.0, -np.asarray(y) * np.asarray(p))))


def mse(y, p) -> float:
    """Mean squared error (util.py:139-141, sklearn ``mean_squared_error``)."""
    y, p = np.asarray(y, dtype=np.float64), np.asarray(p, dtype=np.float64)
    return float(np.mean((y - p) ** 2))


def roc_auc(y, score, pos_label=1) -> float:
    """Area under the ROC curve with the trapezoid rule over *distinct* thresholds, i.e. what
    ``sklearn.metrics.auc(*roc_curve(y, score)[:2])`` returns (ties share one ROC vertex)."""
    y = (np.asarray(y) == pos_label)
    score = np.asarray(score, dtype=np.float64)
    order = np.argsort(-score, kind="mergesort")
    ys, ss = y[order], score[order]
    tp = np.cumsum(ys)
    fp = np.cumsum(~ys)
    last = np.r_[np.flatnonzero(np.diff(ss)), len(ss) - 1]      # last index of each tie block
    tp, fp = np.r_[0, tp[last]], np.r_[0, fp[last]]
    P, N = tp[-1], fp[-1]
    if P == 0 or N == 0:
        return float("nan")
    return float(np.trapezoid(tp / P, fp / N)) if hasattr(np, "trapezoid") else float(np.trapz(tp / P, fp / N))
