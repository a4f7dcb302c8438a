"""Synthetic one-hot data with the *shape* of the reference's real datasets.

The real pipeline (arrange_real_data.py:34-91,145-205,207-253) label-encodes categorical
columns, adds hashed pair features (Amazon), appends a bias column and one-hot encodes, so
every row has a fixed number of ones: one per source column.  No dataset can be downloaded
here, so we draw each categorical column's value from a Zipf-like distribution over that
column's share of the feature space; the last column is the always-on bias.

Shapes (run_approx_coding.sh:26-36): Amazon 26215 x 241915 (9 raw + 34 pairs + bias = 44
ones/row), Covertype 396112 x 15509 (54 + bias = 55), KC-House 17290 x 27654 (18 + bias = 19).
"""
from __future__ import annotations

import numpy as np

from .dataset import Dataset, ell_to_csr

DATASET_SHAPES = {
    "amazon-dataset": dict(n_rows=26215, n_cols=241915, nnz=44, task="logistic"),
    "covtype": dict(n_rows=396112, n_cols=15509, nnz=55, task="logistic"),
    "kc_house_data": dict(n_rows=17290, n_cols=27654, nnz=19, task="linear"),
}


def _column_blocks(n_cols: int, nnz: int, rng) -> np.ndarray:
    """Split the non-bias feature range into ``nnz-1`` categorical blocks of uneven size."""
    k = nnz - 1
    w = rng.dirichlet(np.full(k, 0.6)) * (n_cols - 1 - k)
    sizes = np.floor(w).astype(np.int64) + 1
    sizes[-1] += (n_cols - 1) - sizes.sum()
    return np.concatenate([[0], np.cumsum(sizes)])


def generate_onehot(n_rows: int, n_cols: int, nnz: int, n_partitions: int, seed: int = 0,
                    task: str = "logistic", test_frac: float = 0.25, name: str = "onehot",
                    classes: int = 1) -> Dataset:
    """``task``: logistic (labels +-1), linear (real targets) or softmax (class ids ``0..classes-1`` drawn as
    ``argmax(logits + Gumbel)`` from a hidden ``n_cols x classes`` parameter matrix)."""
    rng = np.random.RandomState(seed)
    rpp = n_rows // n_partitions                      # remainder dropped (arrange_real_data.py:79,85)
    edges = _column_blocks(n_cols, nnz, rng)
    beta_star = rng.standard_normal(n_cols) * 0.5
    W_star = rng.standard_normal((n_cols, classes)) * 0.5 if task == "softmax" else None

    def draw(rows):
        idx = np.empty((rows, nnz), dtype=np.int32)
        for c in range(nnz - 1):
            size = edges[c + 1] - edges[c]
            z = np.minimum(rng.zipf(1.3, rows) - 1, size - 1) if size > 1 else np.zeros(rows, np.int64)
            perm_off = (z * 2654435761 + c * 97) % size      # spread the head over the block
            idx[:, c] = edges[c] + perm_off
        idx[:, nnz - 1] = n_cols - 1                  # bias column (arrange_real_data.py:55)
        if task == "softmax":
            logits = W_star[idx].sum(axis=1)
            return idx, np.argmax(logits + rng.gumbel(size=logits.shape), axis=1).astype(np.float64)
        m = beta_star[idx].sum(axis=1)
        if task == "linear":
            y = m / np.sqrt(nnz) + 0.05 * rng.standard_normal(rows)
        else:
            y = 2.0 * rng.binomial(1, 1.0 / (1.0 + np.exp(-m))) - 1.0
        return idx, y

    parts, ys = [], []
    for _ in range(n_partitions):
        idx, y = draw(rpp)
        parts.append(ell_to_csr(idx, n_cols))
        ys.append(y)
    idx_t, y_t = draw(max(1, int(test_frac * n_rows)))
    return Dataset(parts=parts, y=np.concatenate(ys), X_test=ell_to_csr(idx_t, n_cols), y_test=y_t,
                   n_cols=n_cols, rows_per_part=rpp, name=name)
