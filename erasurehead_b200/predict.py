"""Inference helpers: margins / probabilities / class predictions for trained parameters.

The reference stops at training curves; serving a trained model is one more GEMM-shaped pass,
which on a GPU runs on the tcgen05 evaluation kernel (``csrc/eval_tc.cu``) -- many parameter
vectors (checkpoints, seeds, classes) are scored in one pass over the rows.
"""
from __future__ import annotations

from typing import Union

import numpy as np


def margins(X, betas) -> np.ndarray:
    """``X @ betas.T`` for ``betas`` of shape ``[R, d]`` (or ``[d]``).  ``X`` may be a numpy array, a
    scipy sparse matrix (host path) or a CUDA ``torch.Tensor`` (tensor-core path, returns a tensor)."""
    try:
        import torch
    except Exception:                                   # pragma: no cover
        torch = None
    if torch is not None and isinstance(X, torch.Tensor) and X.is_cuda:
        from .ops import metrics as mops
        B = torch.as_tensor(betas, dtype=torch.float32, device=X.device)
        single = B.dim() == 1
        B = B.reshape(1, -1) if single else B
        d = B.shape[1]
        ld = (d + 31) // 32 * 32
        Xp = X.float()
        if Xp.shape[1] != ld or Xp.stride(1) != 1 or Xp.stride(0) != ld:
            Xn = torch.zeros((X.shape[0], ld), dtype=torch.float32, device=X.device)
            Xn[:, :d] = Xp[:, :d]
            Xp = Xn
        Bp = torch.zeros((B.shape[0], ld), dtype=torch.float32, device=X.device)
        Bp[:, :d] = B
        _, P = mops.eval_tc(Xp, Bp, None, d=d, want_margins=True)
        return P[:, 0] if single else P
    B = np.asarray(betas, dtype=np.float64)
    out = X.dot(B.T if B.ndim == 2 else B)
    return np.asarray(out)


def predict_proba(X, betas):
    """Logistic probabilities ``sigmoid(X beta)``."""
    m = margins(X, betas)
    if isinstance(m, np.ndarray):
        return 1.0 / (1.0 + np.exp(-m))
    return m.sigmoid()


def predict_label(X, betas):
    """Labels in {-1, +1} like the reference's training data."""
    m = margins(X, betas)
    return np.where(m >= 0, 1.0, -1.0) if isinstance(m, np.ndarray) else (m >= 0).float() * 2 - 1


def predict_class(X, W: Union[np.ndarray, "object"]) -> np.ndarray:
    """Arg-max class of a softmax model ``W[d, C]`` (host arrays)."""
    return np.argmax(np.asarray(X.dot(np.asarray(W))), axis=1)
