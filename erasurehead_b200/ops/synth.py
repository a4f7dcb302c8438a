"""Synthetic GMM data generated on the device (generate_data.py:23-45 distribution)."""
from __future__ import annotations

import torch

from ._lib import check, lib, stream_ptr


def pad_cols(d: int, multiple: int = 32) -> int:
    return (d + multiple - 1) // multiple * multiple


def synth_gmm(rows: int, d: int, seed: int, partition: int, task: int = 0, device="cuda", ld: int | None = None):
    """Returns ``(X[rows, ld] (view of the first d columns is the data), y[rows])``; columns
    ``d..ld`` are zero so that the padded row stride is harmless to every kernel."""
    ld = ld or pad_cols(d)
    X = torch.empty((rows, ld), dtype=torch.float32, device=device)
    y = torch.empty(rows, dtype=torch.float32, device=device)
    check(lib().eh_synth_gmm(X.data_ptr(), y.data_ptr(), rows, d, ld, int(seed), int(partition), int(task),
                             stream_ptr()), "eh_synth_gmm")
    return X, y
