"""Evaluation epilogue and result files (reference: every driver's tail, e.g. naive.py:154-209).

For every stored iterate the master evaluates train/test loss (logistic) or MSE (linear) and
the test ROC-AUC, prints one line per iteration and writes ``<prefix>_training_loss.dat``,
``_testing_loss.dat``, ``_auc.dat``, ``_timeset.dat`` (``save_vector`` format) and
``_worker_timeset.dat`` (``save_matrix`` format) under ``results/``.  In addition the new
framework writes ``.npy`` files for the parameter iterates, losses and times (the reference
never persists ``betaset``) and, for linear regression, the files whose writers
are commented out in the reference (Q4).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import numpy as np

from .utils import io, metrics


@dataclass
class Evaluation:
    training_loss: np.ndarray
    testing_loss: np.ndarray
    auc: Optional[np.ndarray]
    accuracy: Optional[np.ndarray] = None      # test accuracy per iterate (classification models; not in the reference)


def evaluate_host(model, data, betaset: np.ndarray, n_eval_parts=None) -> Evaluation:
    """Float64 numpy evaluation of all iterates (one matmul for all rounds)."""
    X_tr, y_tr = data.train_matrix(n_eval_parts)
    X_te, y_te = data.X_test, data.y_test
    R = betaset.shape[0]
    tr, te = np.zeros(R), np.zeros(R)
    auc = np.zeros(R) if model.has_auc else None
    acc = None if model.name == "linear" else np.zeros(R)
    if betaset.ndim == 2:
        P_tr = np.asarray(X_tr.dot(betaset.T))
        P_te = np.asarray(X_te.dot(betaset.T))
        for i in range(R):
            tr[i] = model.loss(y_tr, P_tr[:, i])
            te[i] = model.loss(y_te, P_te[:, i])
            if auc is not None:
                auc[i] = metrics.roc_auc(y_te, P_te[:, i])
            if acc is not None:
                acc[i] = float(np.mean(np.where(P_te[:, i] > 0, 1.0, -1.0) == y_te))
    else:
        for i in range(R):
            tr[i] = model.loss(y_tr, np.asarray(X_tr.dot(betaset[i])))
            M = np.asarray(X_te.dot(betaset[i]))
            te[i] = model.loss(y_te, M)
            if acc is not None:
                acc[i] = float(np.mean(np.argmax(M, axis=1) == y_te.astype(np.int64)))
    return Evaluation(tr, te, auc, acc)


def print_table(ev: Evaluation, timeset: np.ndarray, metric: str = "loss", file=None) -> None:
    """stdout lines of naive.py:198 (logistic) / naive.py:403 (linear)."""
    for i in range(len(timeset)):
        if ev.auc is not None:
            print("Iteration %d: Train Loss = %5.3f, Test Loss = %5.3f, AUC = %5.3f, Total time taken =%5.3f"
                  % (i, ev.training_loss[i], ev.testing_loss[i], ev.auc[i], timeset[i]), file=file)
        else:
            print("Iteration %d: Train Loss = %5.3f, Test Loss = %5.3f, Total time taken =%5.3f"
                  % (i, ev.training_loss[i], ev.testing_loss[i], timeset[i]), file=file)


def write_results(out_dir: str, prefix: str, ev: Optional[Evaluation], timeset, worker_timeset,
                  betaset=None, compat: bool = False, scheme: str = "") -> None:
    io.ensure_dir(out_dir)
    j = lambda suffix: os.path.join(out_dir, f"{prefix}_{suffix}")
    if ev is not None:
        tl_prefix = prefix
        if compat and scheme == "partial_coded":
            # Q3: partial_coded.py:286 writes its training loss under the partialreplication_ prefix
            tl_prefix = prefix.replace("partialcoded_", "partialreplication_")
        io.save_vector(ev.training_loss, os.path.join(out_dir, f"{tl_prefix}_training_loss.dat"))
        io.save_vector(ev.testing_loss, j("testing_loss.dat"))
        if ev.auc is not None:
            io.save_vector(ev.auc, j("auc.dat"))
        np.save(j("training_loss.npy"), ev.training_loss)
        np.save(j("testing_loss.npy"), ev.testing_loss)
        if ev.auc is not None:
            np.save(j("auc.npy"), ev.auc)
        if ev.accuracy is not None and not compat:
            io.save_vector(ev.accuracy, j("accuracy.dat"))
            np.save(j("accuracy.npy"), ev.accuracy)
    io.save_vector(timeset, j("timeset.dat"))
    io.save_matrix(worker_timeset, j("worker_timeset.dat"))
    np.save(j("timeset.npy"), np.asarray(timeset))
    np.save(j("worker_timeset.npy"), np.asarray(worker_timeset))
    if betaset is not None:
        np.save(j("betaset.npy"), np.asarray(betaset))
