"""Synthetic GMM data generated on the device (generate_data.py:23-45 distribution)."""
from __future__ import annotations

import torch

from ._lib import check, lib, stream_ptr


def pad_cols(d: int, multiple: int = 32) -> int:
    return (d + multiple - 1) // multiple * multiple


def synth_gmm(rows: int, d: int, seed: int, partition: int, task: int = 0, device="cuda", ld: int | None = None,
              classes: int = 0):
    """Returns ``(X[rows, ld] (view of the first d columns is the data), y[rows])``; columns
    ``d..ld`` are zero so that the padded row stride is harmless to every kernel."""
    ld = ld or pad_cols(d)
    X = torch.empty((rows, ld), dtype=torch.float32, device=device)
    y = torch.empty(rows, dtype=torch.float32, device=device)
    check(lib().eh_synth_gmm(X.data_ptr(), y.data_ptr(), rows, d, ld, int(seed), int(partition), int(task),
                             int(classes), stream_ptr()), "eh_synth_gmm")
    return X, y


def onehot_edges(d: int, nnz: int, seed: int):
    """Block boundaries of the ``nnz - 1`` categorical source columns over features ``0 .. d-2`` (uneven sizes,
    same rule as the host generator ``data/shapes.py``); int32 numpy array of ``nnz`` entries."""
    import numpy as np
    from ..data.shapes import _column_blocks
    return _column_blocks(d, nnz, np.random.RandomState(seed)).astype(np.int32)


def synth_onehot(rows: int, d: int, nnz: int, seed: int, partition: int, task: int = 0, classes: int = 0,
                 device="cuda", edges=None):
    """One-hot ELL partition of a dataset *shape*, generated on the device: ``(idx[rows, nnz] int32, y[rows])``.
    ``task``: 0 logistic labels +-1, 1 linear targets, 2 class ids ``0..classes-1``."""
    if edges is None:
        edges = torch.from_numpy(onehot_edges(d, nnz, seed)).to(device)
    idx = torch.empty((rows, nnz), dtype=torch.int32, device=device)
    y = torch.empty(rows, dtype=torch.float32, device=device)
    check(lib().eh_synth_onehot(idx.data_ptr(), y.data_ptr(), rows, d, nnz, edges.data_ptr(), int(seed), int(partition),
                                int(task), int(classes), stream_ptr()), "eh_synth_onehot")
    return idx, y
