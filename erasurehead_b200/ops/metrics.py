"""Evaluation operators on the device (reference epilogue naive.py:154-209, util.py:136-141)."""
from __future__ import annotations

import torch

from ._lib import check, lib, stream_ptr


def margins_dense(X: torch.Tensor, Bt: torch.Tensor, d: int | None = None) -> torch.Tensor:
    """``P[rows, R] = X[:, :d] @ Bt[:, :d].T`` for all ``R`` stored iterates at once."""
    assert X.dtype == torch.float32 and Bt.dtype == torch.float32 and X.stride(1) == 1 and Bt.stride(1) == 1
    d = int(d or Bt.shape[1])
    R = Bt.shape[0]
    P = torch.empty((X.shape[0], R), dtype=torch.float32, device=X.device)
    check(lib().eh_margins_dense(X.data_ptr(), X.shape[0], d, X.stride(0), Bt.data_ptr(), R, Bt.stride(0),
                                 P.data_ptr(), stream_ptr()), "eh_margins_dense")
    return P


def margins_ell(idx: torch.Tensor, Bt: torch.Tensor) -> torch.Tensor:
    assert idx.dtype == torch.int32 and idx.is_contiguous() and Bt.dtype == torch.float32
    R = Bt.shape[0]
    P = torch.empty((idx.shape[0], R), dtype=torch.float32, device=idx.device)
    check(lib().eh_margins_ell(idx.data_ptr(), idx.shape[0], idx.shape[1], Bt.data_ptr(), R, Bt.stride(0),
                               P.data_ptr(), stream_ptr()), "eh_margins_ell")
    return P


def loss_columns(P: torch.Tensor, y: torch.Tensor, kind: int) -> torch.Tensor:
    """Mean loss per column of ``P`` (kind 0: logistic with labels +-1; 1: squared error)."""
    assert P.is_contiguous() and P.dtype == torch.float32 and y.dtype == torch.float32
    out = torch.zeros(P.shape[1], dtype=torch.float64, device=P.device)
    check(lib().eh_loss_columns(P.data_ptr(), y.data_ptr(), P.shape[0], P.shape[1], int(kind), out.data_ptr(),
                                stream_ptr()), "eh_loss_columns")
    return out / P.shape[0]


def auc(scores: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """Tie-aware ROC-AUC of one score vector (labels +1/-1) -> 0-d float64 tensor."""
    scores = scores.contiguous().float()
    labels = labels.contiguous().float()
    n = scores.numel()
    nbytes = lib().eh_auc_scratch_bytes(n)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=scores.device)
    out = torch.empty((), dtype=torch.float64, device=scores.device)
    check(lib().eh_auc(scores.data_ptr(), labels.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), nbytes,
                       stream_ptr()), "eh_auc")
    return out


def auc_batch(scores: torch.Tensor, labels: torch.Tensor, max_elems: int = 1 << 27) -> torch.Tensor:
    """Tie-aware ROC-AUC of every row of ``scores[R, n]`` (one row per stored iterate) -> float64 ``[R]``.
    One sort + two scans for all rows (``eh_auc_batch``); very large inputs go in chunks of rows."""
    scores = scores.contiguous().float()
    labels = labels.contiguous().float()
    R, n = scores.shape
    out = torch.empty(R, dtype=torch.float64, device=scores.device)
    per = max(1, min(R, max_elems // max(1, n)))
    nbytes = lib().eh_auc_batch_scratch_bytes(n, per)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=scores.device)
    for r0 in range(0, R, per):
        k = min(per, R - r0)
        check(lib().eh_auc_batch(scores[r0:r0 + k].data_ptr(), labels.data_ptr(), n, k, out[r0:r0 + k].data_ptr(),
                                 scratch.data_ptr(), nbytes, stream_ptr()), "eh_auc_batch")
    return out


def eval_tc(X: torch.Tensor, Bt: torch.Tensor, y: torch.Tensor | None = None, kind: int = 0, d: int | None = None,
            want_margins: bool = False, loss_sums: torch.Tensor | None = None):
    """tcgen05/TMEM evaluation: one pass over ``X`` for all ``R`` iterates (TF32 operands).

    Returns ``(loss_sums[R] float64 or None, P[rows, R] float32 or None)``.  ``loss_sums`` holds
    column SUMS (accumulates into the tensor passed in, so partitions can be chained)."""
    assert X.dtype == torch.float32 and Bt.dtype == torch.float32 and X.stride(1) == 1 and Bt.stride(1) == 1
    d = int(d or Bt.shape[1])
    R = Bt.shape[0]
    if y is not None and loss_sums is None:
        loss_sums = torch.zeros(R, dtype=torch.float64, device=X.device)
    P = torch.empty((X.shape[0], R), dtype=torch.float32, device=X.device) if want_margins else None
    check(lib().eh_eval_tc(X.data_ptr(), X.shape[0], d, X.stride(0), Bt.data_ptr(), R, Bt.stride(0),
                           0 if y is None else y.data_ptr(), int(kind),
                           0 if (y is None or loss_sums is None) else loss_sums.data_ptr(),
                           0 if P is None else P.data_ptr(), R, stream_ptr()), "eh_eval_tc")
    return (loss_sums if y is not None else None), P
