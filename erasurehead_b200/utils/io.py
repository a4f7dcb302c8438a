"""File formats of the reference, byte-compatible.

* dense partition ``k.dat``: one row per line, space-separated ``str(float)`` (util.py:26-30),
  read back with ``np.loadtxt`` (util.py:13-15);
* sparse partition ``k.npz``: CSR arrays under ``data, indices, indptr, shape`` (util.py:17-24);
* vectors (labels, per-round losses/times): one ``"%5.3f "`` per line (util.py:32-36).
"""
from __future__ import annotations

import os

import numpy as np
from scipy.sparse import csr_matrix

__all__ = ["load_data", "save_matrix", "save_vector", "save_sparse_csr", "load_sparse_csr",
           "ensure_dir"]


def ensure_dir(path: str) -> str:
    os.makedirs(path, exist_ok=True)
    return path


def load_data(path: str) -> np.ndarray:
    return np.loadtxt(path, dtype=float)


def save_matrix(m, path: str) -> None:
    with open(path, "w") as f:
        for row in m:
            f.write(" ".join(repr(float(x)) for x in row) + "\n")


def save_vector(v, path: str) -> None:
    with open(path, "w") as f:
        for x in v:
            f.write("%5.3f" % x + " \n")


def save_sparse_csr(path: str, array) -> None:
    array = csr_matrix(array)
    np.savez(path, data=array.data, indices=array.indices, indptr=array.indptr, shape=array.shape)


def load_sparse_csr(path: str) -> csr_matrix:
    loader = np.load(path if path.endswith(".npz") else path + ".npz")
    return csr_matrix((loader["data"], loader["indices"], loader["indptr"]), shape=loader["shape"])
